"""C5 of BASELINE.json at the size SURVEY.md 8(d) names (65 536 x 65 536): N sweep over a random uniform (1 %) and a banded matrix, every
kernel that accepts the chain and the width -- where row blocking, column-direction blocking, nnz blocking, MERGE PATH, the B200 gather
kernels and the tensor-core kernels cross over, and what the library's own choice (GS_KERNEL_AUTO) costs.
Writes a markdown table (us per launch, best kernel per cell in bold) to gpurun_out/r2_c5_sweep.md.  Every timed plan is first checked
against the fp64 oracle on a row sample (fp32: 1e-5, HALF: 1e-2 of |ref|, floor 1e-6 / 1e-2 of the row's L1)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth

M = int(os.environ.get("C5_ROWS", 65536))
NS = [int(x) for x in os.environ.get("C5_NS", "1,2,4,8,16,32,64,128,256,512").split(",")]
MP_T = [("merge_path_thread_operator", 64), ("thread_total_reduce_operator", 0, 1, 1), ("grid_block_operator", 64, 8, 16, 1)]
MP_W = [("merge_path_warp_operator", 1024), ("warp_total_reduce_operator", 1), ("grid_block_operator", 64, 32, 4, 1)]
# chain -> (how to build it, sparse coarsen, [(label, kernel, half)])
CHAINS = [
    ("thread_total", "thread_total", 8, [("thread_total (reference-shaped) f32", None, False), ("b200_row_smem f32", "b200_row_smem", False), ("b200_row_smem f16", "b200_row_smem", True),
                                          ("b200_merge on the row chain f32", "b200_merge", False), ("b200_tcgen05 f16", "b200_tcgen05", True), ("b200_sp24 f16", "b200_sp24", True),
                                          ("AUTO f32", "auto", False), ("AUTO f16", "auto", True)]),
    ("warp_total", "warp_total", 1, [("warp_total (reference-shaped) f32", None, False)]),
    ("thread_bit_map", "thread_bit_map", 1, [("thread_bit_map (reference-shaped, nnz direction) f32", None, False), ("b200_nnz f32", "b200_nnz", False)]),
    ("warp_segment", "warp_segment", 1, [("warp_segment (reference-shaped, nnz direction) f32", None, False)]),
    ("warp_bitmap", "warp_bitmap", 1, [("warp_bitmap (reference-shaped, col direction) f32", None, False), ("b200_merge on the col-direction chain f32", "b200_merge", False)]),
    ("merge_path_thread(64)", MP_T, 1, [("b200_merge on merge_path_thread(64) f32", None, False), ("b200_merge on merge_path_thread(64) f16", None, True)]),
    ("merge_path_warp(1024)", MP_W, 1, [("b200_merge on merge_path_warp(1024) f32", None, False)]),
]
WEIGHT_SIZED = M * M * 1.125 <= 2 ** 31
eng = api.Engine(0)
out = []
t_start = time.time()
for mat, kind, kw in (("uniform 1 %", "uniform", dict(density=0.01)), ("banded, half bandwidth 64", "banded", dict(half_bandwidth=64))):
    r, c = synth.synth_matrix(kind, M, M, seed=5, **kw)
    vals = synth.synth_values(len(r), seed=5).astype(np.float16).astype(np.float32)
    indptr = np.concatenate([[0], np.cumsum(np.bincount(r, minlength=M))]).astype(np.int64)
    sample = np.unique(np.random.default_rng(1).integers(0, M, 256))
    table, picked, alg = {}, {}, {}
    labels = [lab for _, _, _, ks in CHAINS for lab, _, _ in ks]
    for N in NS:
        B = synth.synth_dense(M, N, seed=6).astype(np.float16).astype(np.float32)
        for cname, build, S, kernels in CHAINS:
            try:
                ms = api.MetaDataSet.from_coo(M, M, r, c, vals)
                s = api.Session(ms)
                if isinstance(build, str):
                    s.run_family(build, N, S, 1)
                else:
                    api.config_set("DENSE_MATRIX_SIZE", N)
                    api.config_set("VECTOR_WIDTH", min(N, 32))
                    for op in build:
                        s.add_and_run(op[0], *op[1:])
                    s.compile()
            except api.GsError as e:
                for lab, _, _ in kernels:
                    table[(lab, N)] = None
                continue
            uploaded = {}
            for lab, kernel, half in kernels:
                if os.environ.get("C5_ONLY") and os.environ["C5_ONLY"] not in lab:
                    table[(lab, N)] = None
                    continue
                if kernel in ("b200_tcgen05", "b200_sp24") and not WEIGHT_SIZED:
                    table[(lab, N)] = None  # the tensor-core layouts are built for weight-sized matrices (4.8 GB of image here)
                    continue
                try:
                    if half not in uploaded:
                        eng.upload_session(s, half=half)
                        uploaded = {half: True}
                    plan = eng.plan(s, N, half=half, kernel=kernel)
                    plan.set_dense(B)
                    plan.run(1)
                    C = plan.output().astype(np.float64)
                    ok = True
                    for i in sample:
                        cols_i, v_i = c[indptr[i]:indptr[i + 1]], vals[indptr[i]:indptr[i + 1]].astype(np.float64)
                        ref = v_i @ B[cols_i].astype(np.float64)
                        floor = np.abs(v_i) @ np.abs(B[cols_i].astype(np.float64))
                        tol = 1e-2 if half else 1e-5
                        ok &= bool((np.abs(C[i] - ref) <= tol * np.abs(ref) + (1e-2 if half else 1e-6) * floor + 2.0 ** -25).all())
                    ms1 = plan.run(3)
                    reps = int(max(5, min(200, 200.0 / max(ms1, 1e-3))))  # ~0.2 s of launches
                    us = plan.run(reps) * 1e3
                    table[(lab, N)] = (us, ok)
                    alg[(lab, N)] = plan.algorithmic_bytes
                    if kernel == "auto":
                        picked[(lab, N)] = plan.kernel_name.split("<")[0]
                    del plan
                except api.GsError as e:
                    table[(lab, N)] = None
        print("N=%d done at %.0f s" % (N, time.time() - t_start), flush=True)
    out.append("### %s, %d x %d, %d nnz (us per launch; -- = the kernel does not take this chain / width / size)\n" % (mat, M, M, len(r)))
    out.append("| chain -> kernel | " + " | ".join("N=%d" % n for n in NS) + " |")
    out.append("|---|" + "---|" * len(NS))
    best = {n: min((table[(k, n)][0] for k in labels if table.get((k, n)) and "AUTO" not in k), default=None) for n in NS}
    for lab in labels:
        cells = []
        for n in NS:
            t = table.get((lab, n))
            if t is None:
                cells.append("--")
            else:
                us, ok = t
                txt = "%.1f" % us + ("" if ok else " (PARITY FAIL)") + (" [%s]" % picked[(lab, n)].replace("k_b200_", "") if (lab, n) in picked else "")
                cells.append("**%s**" % txt if best[n] is not None and us == best[n] else txt)
        out.append("| %s | %s |" % (lab, " | ".join(cells)))
    out.append("")
    bl = {n: [k for k in labels if table.get((k, n)) and table[(k, n)][0] == best[n]][0] for n in NS if best[n]}
    out.append("Best kernel per width: " + ", ".join("N=%d %.0f GFLOP/s = %.0f %% of the HBM roofline on SURVEY 8(d) bytes (%s)" % (
        n, 2.0 * len(r) * n / (best[n] * 1e-6) / 1e9, 100.0 * alg[(bl[n], n)] / (best[n] * 1e-6) / 6470.2e9, bl[n]) for n in NS if best[n]))
    out.append("")
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/r2_c5_sweep.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
