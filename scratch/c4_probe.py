import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth
from oracle.oracle_check import csr_from_coo, oracle_spmm_f64acc, oracle_spmm_absrow
rows = int(os.environ.get("ROWS", 1000000)); N = 32
t=time.time()
r, c = synth.synth_matrix("powerlaw", rows, rows, seed=4, avg_deg=20.0, alpha=2.1, cap=100000)
vals = synth.synth_values(len(r), seed=4)
B = synth.synth_dense(rows, N, seed=4)
print("gen %.1fs nnz %d maxrow %d" % (time.time()-t, len(r), np.bincount(r).max()), flush=True)
eng = api.Engine(0)
indptr = csr_from_coo(rows, r)
sub = slice(0, 20000)
ref = None
for family, S, Cc, kernel in (("thread_total", 1, 1, None), ("thread_total", 4, 1, None), ("thread_total", 8, 1, "b200_row_smem"), ("thread_total", 1, 1, "b200_row_smem"), ("warp_total", 1, 1, None), ("warp_total", 1, 1, "b200_row_smem"),
                              ("warp_segment", 1, 1, None), ("thread_bit_map", 1, 1, None), ("warp_bitmap", 1, 1, None)):
    try:
        t=time.time()
        ms = api.MetaDataSet.from_coo(rows, rows, r, c, vals)
        s = api.Session(ms).run_family(family, N, S, Cc)
        th = time.time()-t
        eng.upload_session(s)
        plan = eng.plan(s, N, kernel=kernel)
        plan.set_dense(B)
        plan.run(1)
        C = plan.output()
        if ref is None:
            ip = indptr[:20001]
            ref = oracle_spmm_f64acc(20000, N, ip, c[:ip[-1]], vals[:ip[-1]], B)
        err = np.abs(C[sub] - ref).max() / np.abs(ref).max()
        m = plan.run(10)
        print("%-14s S=%d kernel=%-14s host %.1fs  %.3f ms  alg %.0f GB/s  %.1f GFLOP/s  err %.2g  %s" % (family, S, kernel, th, m, plan.algorithmic_bytes / m / 1e6, 2.0 * len(r) * N / m / 1e6, err, plan.kernel_name), flush=True)
        del plan
    except Exception as e:
        print(family, S, kernel, "FAILED", str(e)[:150], flush=True)
