"""C5 of BASELINE.json: N sweep over a random uniform and a banded matrix, every kernel family that accepts the chain and
the width -- where row blocking, nnz blocking, the B200 gather kernels and the tcgen05 tile kernel cross over.
Writes a markdown table (µs per launch, best kernel per cell in bold) to gpurun_out/c5_sweep.md.  Every timed plan is
first checked against the fp64 oracle on a row sample (fp32: 1e-5, HALF: 1e-2)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth
from oracle import oracle_check as oc

M = int(os.environ.get("C5_ROWS", 16384))
NS = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512]
# (label, chain, sparse coarsen, kernel override, half)
KERNELS = [("thread_total f32", "thread_total", 4, None, False), ("warp_total f32", "warp_total", 1, None, False),
           ("thread_bit_map f32", "thread_bit_map", 1, None, False), ("warp_segment f32", "warp_segment", 1, None, False),
           ("b200_nnz f32", "thread_bit_map", 1, "b200_nnz", False),
           ("b200_row_smem f16", "thread_total", 8, "b200_row_smem", True), ("b200_tcgen05 f16", "thread_total", 8, "b200_tcgen05", True),
           ("b200_sp24 f16", "thread_total", 8, "b200_sp24", True),
           ("thread_total f16", "thread_total", 8, None, True)]
eng = api.Engine(0)
out = []
for mat, kind, kw in (("uniform 1 %", "uniform", dict(density=0.01)), ("banded, half bandwidth 64", "banded", dict(half_bandwidth=64))):
    r, c = synth.synth_matrix(kind, M, M, seed=5, **kw)
    vals = synth.synth_values(len(r), seed=5).astype(np.float16).astype(np.float32)
    indptr = oc.csr_from_coo(M, r)
    sample = np.unique(np.random.default_rng(1).integers(0, M, 512))
    table = {}
    for label, chain, S, kernel, half in KERNELS:
        for N in NS:
            B = synth.synth_dense(M, N, seed=6).astype(np.float16).astype(np.float32)
            try:
                ms = api.MetaDataSet.from_coo(M, M, r, c, vals)
                s = api.Session(ms).run_family(chain, N, S, 1)
                eng.upload_session(s, half=half)
                plan = eng.plan(s, N, half=half, kernel=kernel)
                plan.set_dense(B)
                plan.run(1)
                C = plan.output().astype(np.float64)
                # parity on a row sample
                ok = True
                for i in sample:
                    cols_i, v_i = c[indptr[i]:indptr[i + 1]], vals[indptr[i]:indptr[i + 1]].astype(np.float64)
                    ref = v_i @ B[cols_i].astype(np.float64)
                    floor = np.abs(v_i) @ np.abs(B[cols_i].astype(np.float64))
                    tol = 1e-2 if half else 1e-5
                    ok &= bool((np.abs(C[i] - ref) <= tol * np.abs(ref) + tol * floor + 2.0 ** -25).all())
                plan.run(20)
                us = plan.run(200) * 1e3
                table[(label, N)] = (us, ok)
            except api.GsError as e:
                table[(label, N)] = None
    out.append("### %s, %d x %d, %d nnz (µs per launch; — = the kernel does not take this chain / width)\n" % (mat, M, M, len(r)))
    out.append("| kernel | " + " | ".join("N=%d" % n for n in NS) + " |")
    out.append("|---|" + "---|" * len(NS))
    best = {n: min((table[(k[0], n)][0] for k in KERNELS if table.get((k[0], n))), default=None) for n in NS}
    for label, *_ in KERNELS:
        cells = []
        for n in NS:
            t = table.get((label, n))
            if t is None:
                cells.append("—")
            else:
                us, ok = t
                txt = "%.1f" % us + ("" if ok else " (PARITY FAIL)")
                cells.append("**%s**" % txt if best[n] is not None and us == best[n] else txt)
        out.append("| %s | %s |" % (label, " | ".join(cells)))
    out.append("")
    out.append("GFLOP/s of the best kernel: " + ", ".join("N=%d %.0f" % (n, 2.0 * len(r) * n / (best[n] * 1e-6) / 1e9) for n in NS if best[n]))
    out.append("")
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/c5_sweep.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
