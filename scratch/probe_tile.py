"""One-tile probe of the tcgen05 path: which (row, k, n) each output element actually read."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth
from oracle.oracle_check import csr_from_coo, oracle_spmm_f64acc

eng = api.Engine(0)
for N in (32, 16, 64, 128):
    rows, cols = 128, 64
    r = np.arange(rows, dtype=np.int64)
    c = (r * 5 + 3) % 64
    vals = np.ones(rows, np.float32)
    B = (np.arange(cols)[:, None] * 32 + (np.arange(N)[None, :] % 32)).astype(np.float32)  # <= 2047: exact in fp16
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("warp_total", N, 1, 1)
    eng.upload_session(s, half=True)
    plan = eng.plan(s, N, half=True, kernel="b200_tcgen05")
    plan.set_dense(B)
    plan.run(1)
    C = plan.output().astype(np.float32)
    want = B[c]
    ok = np.array_equal(C, want)
    print("N=%d one-tile probe ok=%s kernel=%s" % (N, ok, plan.kernel_name))
    if not ok:
        bad = np.argwhere(C != want)
        print("  mismatches:", len(bad), "of", C.size)
        for i in list(range(0, 16)) + [31, 32, 63, 64, 127]:
            got = C[i]
            print("  row %3d want k=%2d got k=%s n=%s" % (i, c[i], (got // 32).astype(int)[:8].tolist(), (got % 32).astype(int)[:8].tolist()))
    del plan

for (rows, cols, N, dens) in ((1000, 4096, 32, 0.1), (300, 1000, 16, 0.05), (11008, 4096, 32, 0.3)):
    r, c = synth.synth_matrix("uniform", rows, cols, seed=7, density=dens)
    vals = synth.synth_values(len(r), seed=7).astype(np.float16).astype(np.float32)
    B = synth.synth_dense(cols, N, seed=7).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    eng.upload_session(s, half=True)
    plan = eng.plan(s, N, half=True, kernel="b200_tcgen05")
    plan.set_dense(B)
    plan.run(1)
    C = plan.output().astype(np.float64)
    ref = oracle_spmm_f64acc(rows, N, csr_from_coo(rows, r), c, vals, B)
    err = np.abs(C - ref).max() / np.abs(ref).max()
    ms_it = plan.run(200)
    print("%dx%d N=%d dens=%.2f max err/max|ref| = %.3g   %.2f us/launch  bytes %d  -> %.0f GB/s  %s" % (rows, cols, N, dens, err, ms_it * 1e3, plan.algorithmic_bytes,
          plan.algorithmic_bytes / (ms_it * 1e-3) / 1e9, plan.kernel_name))
    del plan
