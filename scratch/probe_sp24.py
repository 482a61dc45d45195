"""First contact of the 2:4 kernel with the hardware: (1) product of a 128 x 128 2:4 matrix against the oracle,
(2) the metadata columns as tensor memory holds them, (3) where single elements land: (r, k) -> (row, k') seen in C."""
import os
import sys

os.environ["GS_SP24_PROBE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth

eng = api.Engine(0)
rows, cols, N = 128, 128, 16


def run(r, c, vals, B, chain="warp_total"):
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family(chain, N, 1, 1)
    eng.upload_session(s, half=True)
    plan = eng.plan(s, N, half=True, kernel="b200_sp24")
    plan.set_dense(B)
    plan.run(1)
    return plan.output().astype(np.float32), plan


r, c = synth.synth_matrix("nm24", rows, cols, seed=95)
vals = synth.synth_values(len(r), seed=1).astype(np.float16).astype(np.float32)
B = synth.synth_dense(cols, N, seed=2).astype(np.float16).astype(np.float32)
A = np.zeros((rows, cols), np.float64)
A[r, c] = vals
C, plan = run(r, c, vals, B)
print(plan.kernel_name)
ref = A @ B.astype(np.float64)
print("product: max abs err %.3g (max |ref| %.3g)" % (np.abs(C - ref).max(), np.abs(ref).max()))
got = plan.sp24_probe()
keep = A != 0
want = np.zeros((128, 4), dtype=np.uint32)
for L in range(128):
    for j in range(4):
        w = 0
        for half in range(2):
            row = (L & 7) + 16 * (L >> 4) + 8 * half
            k0 = 32 * j + 16 * ((L >> 3) & 1)
            for q in range(4):
                nz = np.nonzero(keep[row, k0 + 4 * q:k0 + 4 * q + 4])[0]
                i0, i1 = (int(nz[0]), int(nz[1])) if len(nz) == 2 else (0, 1)
                w |= (i0 | (i1 << 2)) << (16 * half + 4 * q)
        want[L, j] = w
print("metadata in tensor memory: %d of 512 words differ" % int((got != want).sum()))
if (got != want).any():
    for L in range(0, 128, 9):
        print("  lane %3d got %s want %s" % (L, " ".join("%08x" % x for x in got[L]), " ".join("%08x" % x for x in want[L])))
# single elements
Bk = np.zeros((cols, N), np.float32)
Bk[:, 0] = np.arange(cols) + 1
Bk[:, 1] = 1
bad = 0
for (rr, kk) in [(0, 0), (0, 1), (0, 2), (0, 3), (1, 0), (7, 5), (8, 0), (8, 3), (9, 17), (15, 31), (16, 32), (17, 47), (31, 63), (32, 64), (64, 96), (100, 101), (127, 127), (127, 0), (5, 126)]:
    C, _ = run(np.array([rr]), np.array([kk]), np.array([1.0], np.float32), Bk)
    nz = np.argwhere(C[:, :2] != 0)
    seen = [(int(i), float(C[i, 0]) - 1, float(C[i, 1])) for i in sorted(set(nz[:, 0]))]
    ok = seen == [(rr, float(kk), 1.0)]
    bad += not ok
    print("  element (%3d, %3d) -> rows/k'/weight seen: %s %s" % (rr, kk, seen, "" if ok else "<-- WRONG"))
print("single elements: %d wrong" % bad)
