"""plan creation (the kernel's own layout built from the chain's device arrays) per kernel, wall time"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth
eng = api.Engine(0)
def t_plan(desc, N, half, kernel, reps=2):
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        p = eng.plan(desc, N, half=half, kernel=kernel)
        best = min(best, time.perf_counter() - t0)
        name = p.kernel_name
        del p
    return best * 1e3, name
rows, cols = 11008, 4096
r, c = synth.synth_matrix("uniform", rows, cols, seed=2, density=0.30)
v = synth.synth_values(len(r), seed=2).astype(np.float16).astype(np.float32)
for N, half, kernels in ((8, True, ("b200_sp24", "b200_row_smem", "b200_merge", None)), (32, True, ("b200_sp24", "b200_tcgen05", "b200_merge"))):
    d, ms = eng.transform_coo("thread_total", rows, cols, r, c, v, N, 8, 1, half=half, as_id=1)
    for k in kernels:
        ms_p, name = t_plan(d, N, half, k)
        print("C2 N=%d %-14s plan %.1f ms  %s" % (N, k, ms_p, name[:70]))
rows, cols, r, c = synth.c1_standin()
for N in (8,):
    d, ms = eng.transform_coo("thread_total", rows, cols, r, c, None, N, 4, 1, as_id=2)
    for k in ("b200_merge", "b200_row_smem", "auto"):
        ms_p, name = t_plan(d, N, False, k)
        print("C1 N=%d %-14s plan %.1f ms  %s" % (N, k, ms_p, name[:70]))
rows = 1_000_000
r, c = synth.synth_matrix("powerlaw", rows, rows, seed=4, avg_deg=20.0, alpha=2.1)
d, ms = eng.transform_coo("thread_bit_map", rows, rows, r, c, None, 32, 1, 1, as_id=3)
print("C4/10 chain on device %.1f ms" % ms)
for k in ("b200_nnz", "b200_merge", None):
    ms_p, name = t_plan(d, 32, False, k)
    print("C4/10 N=32 %-14s plan %.1f ms  %s" % (k, ms_p, name[:70]))
