#!/usr/bin/env python3
"""bench.py -- SpMM GFLOP/s (2*nnz*N/t) and % of the HBM roofline on B200.

Workload (BASELINE.json configs[1]): synthetic pruned Llama-2-7B FFN weight 11008 x 4096, 70 %
unstructured sparsity, HALF storage with fp32 accumulation, N dense columns (default 8; --n 16/32).
A step = one SpMM C = A*B through the reference's thread_total operator chain (sort_operator ->
row blocking with sparse-coarsen-8 padding -> thread_total_reduce) executed by the B200 row kernel (N=8) or the
B200 tile kernel (sparse tile decompress to shared memory + tcgen05, N >= 16).
Successive steps use successive device copies of A (R copies, R*bytes(A) > L2) so every launch
streams A from HBM.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
  python bench.py --impl reference ...                     (the reference's own CPU SpMM on the host cores)

Prints ONE JSON line (rank 0).  `value` = whole-job GFLOP/s with inputs resident in HBM; `e2e` = the
same metric through gs_spmm_host (pinned host B -> H2D -> kernel -> D2H C inside the timed region).
Multi-GPU: weak scaling -- the global matrix is G FFN-sized row blocks stacked (G*11008 x 4096), cut with
the reference's nnz-balanced row rule, one block per rank; B replicated, C rows disjoint, no collective
on the data path.  The same line carries a `c4` block at every N (--no-c4 skips it): BASELINE.json configs[3], ONE fixed
10 M-row power-law matrix cut into N nnz-balanced row blocks (strong scaling; efficiency = c4.ms_per_step at N=1 divided by
N x c4.ms_per_step at N), with all-rank parity, clocks, and the optional NCCL all-gather of C timed beside it.
A failed parity gate nulls `value` and exits 1.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS, DENSITY = 11008, 4096, 0.30
ROWS = int(os.environ.get("GS_BENCH_ROWS", ROWS))  # experiments only; the reported workload is 11008 rows


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--dense-n", dest="n", type=int, default=8,
                    help="dense width N (8 = the bench line; 16..256 run the tile kernel). Under torchrun spell it --dense-n (its parser claims --n)")
    ap.add_argument("--density", type=float, default=0.30, help="c2: keep probability of the pruned weight (0.30 = C2, 0.10 = C3)")
    ap.add_argument("--kernel", default="auto", help="auto: the library picks (GS_KERNEL_AUTO); or b200_row_smem | b200_tcgen05 | b200_sp24")
    ap.add_argument("--pattern", default="unstructured", choices=["unstructured", "nm24"],
                    help="c2: Bernoulli(density) mask, or exactly 2 of every 4 consecutive columns kept (C3's 2:4 half; runs b200_sp24 under --kernel auto)")
    ap.add_argument("--copies", type=int, default=0, help="device copies of A to rotate over (0 = enough to exceed L2 3x)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4"], help="c2: pruned FFN weight (the bench line); c4: power-law matrix, fp32, N=32, "
                    "one fixed matrix cut into nnz-balanced row blocks over the ranks (strong scaling)")
    ap.add_argument("--rows", type=int, default=10_000_000, help="c4: rows = cols of the power-law matrix")
    ap.add_argument("--avg-deg", type=float, default=20.0, help="c4: mean non-zeros per row")
    ap.add_argument("--family", default="thread_bit_map", help="c4: operator chain / kernel family")
    ap.add_argument("--no-yardstick", action="store_true", help="skip the cuSPARSE yardstick (N=1 GPU only)")
    ap.add_argument("--no-c4", action="store_true", help="c2: leave the c4 block (strong scaling of the power-law matrix) out of the line")
    ap.add_argument("--c4-steps", type=int, default=20)
    ap.add_argument("--c4-kernel", default="auto", help="c4: auto = b200_nnz on the bit-map chains, reference = the chain's own family kernel, or a kernel name")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def traffic_bytes(N, half, world, args):
    """DRAM bytes per launch from the committed ncu --set full capture of this workload (profiles/traffic.json), else None"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if args == "c4":
        return json.load(open(p)).get("C4_N32_f32_nnz", {}).get("bytes") if os.path.exists(p) else None
    if not (os.path.exists(p) and half and ROWS == 11008):
        return None
    t = json.load(open(p))
    if args.pattern == "nm24":
        key = "C3_24_N%d_half_sp24" % N if args.kfam == "b200_sp24" else None
    elif abs(DENSITY - 0.30) > 1e-9:
        key = None
    else:
        key = {"b200_row_smem": "C2_N%d_half_thread_total_S8", "b200_tcgen05": "C2_N%d_half_thread_total_S8_tile", "b200_sp24": "C2_N%d_half_sp24"}.get(args.kfam)
        key = key % N if key else None
    return t.get(key, {}).get("bytes") if key else None


PATTERN = "unstructured"


def make_weight(seed):
    from generalsparse_b200 import synth
    if PATTERN == "nm24":
        r, c = synth.synth_matrix("nm24", ROWS, COLS, seed=seed)
    else:
        r, c = synth.synth_matrix("uniform", ROWS, COLS, seed=seed, density=DENSITY)
    v = synth.synth_values(len(r), seed=seed).astype(np.float16).astype(np.float32)
    return r, c, v


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        inside = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.06] or [r for (_, r) in self.rows]
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no sample"]}
        sm = sorted(float(r[0]) for r in inside)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in inside)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(inside[0][1]), "reasons": reasons, "samples": len(inside)}


def device_transform_block(*a, **kw):
    """the block below, or what went wrong with it: a side measurement must not take the bench line down"""
    try:
        return _device_transform_block(*a, **kw)
    except Exception as e:  # noqa: BLE001
        os.environ.pop("GS_SP24_DEVICE_BUILD", None)
        return {"error": "%s: %s" % (type(e).__name__, e)}


def _device_transform_block(eng, sess, family, rows, cols, r, c, v, N, S, half, host_id, host_pipeline_s, plan_kernel="auto"):
    """the chain's transform steps as device passes (gs_transform_host_coo) beside the host pipeline that built the timed plan:
    every array the chain registers must come out byte-equal; device_ms = CUDA events around the passes alone, wall_ms adds the H2D of the COO."""
    eng.transform_coo(family, rows, cols, r, c, v, N, S, 1, half=half, as_id=900)  # warm-up (allocations, module load)
    dev_ms, wall = 1e30, 1e30
    for _ in range(3):  # best of three: the passes are a dozen small kernels with host round trips between them, a busy host core shows up at once
        t0 = time.perf_counter()
        dd, ms1 = eng.transform_coo(family, rows, cols, r, c, v, N, S, 1, half=half, as_id=900)
        wall, dev_ms = min(wall, time.perf_counter() - t0), min(dev_ms, ms1)
    keys = [k[: k.rfind("_")] for k in sess.deps()]
    same = True
    for k in keys:
        a, b = eng.download_array("%s_%d" % (k, host_id)), eng.download_array(k + "_900")
        same = same and a.dtype == b.dtype and a.shape == b.shape and bool(np.array_equal(a.view(np.uint8), b.view(np.uint8)))
    hd = sess.plan_desc()
    same = same and (hd.stored_nnz, tuple(hd.grid), tuple(hd.block), hd.vector_width) == (dd.stored_nnz, tuple(dd.grid), tuple(dd.block), dd.vector_width)
    # plan creation on the device-built chain: the kernel's own layout (the 2:4 image: csrc/device/sp24_build.cu) built by device passes, against
    # the host builder (download of the chain's arrays, host build, upload) on the same arrays
    plan_s = {}
    for mode in ("0", "1", "1"):
        os.environ["GS_SP24_DEVICE_BUILD"] = mode
        t0 = time.perf_counter()
        pp = eng.plan(dd, N, half=half, kernel=plan_kernel)
        plan_s[mode] = time.perf_counter() - t0
        pname = pp.kernel_name
        del pp
    os.environ.pop("GS_SP24_DEVICE_BUILD", None)
    return {"plan_build_ms": plan_s["1"] * 1e3, "plan_build_host_builder_ms": plan_s["0"] * 1e3, "plan_kernel": pname.split("<")[0],"device_ms": dev_ms, "wall_ms_with_h2d_of_coo": wall * 1e3, "host_pipeline_s": host_pipeline_s, "arrays": len(keys), "identical_to_host_pipeline": bool(same),
            "speedup_vs_host_pipeline": host_pipeline_s * 1e3 / max(dev_ms, 1e-6)}


def cusparse_yardstick(rows, cols, r, c, v, B_t, steps=200):
    """the one external sanity bar the reference ships -- baseline/base_cusparse/spmm.cu:82-162, cusparseSpMM on CSR with
    10 warm-ups, gflops = 2 * nnz * N / t -- here through torch's CSR matmul (cuSPARSE SpMM underneath; a library call, comparison only, never
    on the product path).  B_t: K x N torch tensor on the device; values take its dtype."""
    import torch
    try:
        crow = torch.from_numpy(np.concatenate([[0], np.cumsum(np.bincount(r, minlength=rows))]).astype(np.int32)).cuda()
        A = torch.sparse_csr_tensor(crow, torch.from_numpy(c.astype(np.int32)).cuda(), torch.from_numpy(v).to(B_t.dtype).cuda(), size=(rows, cols))
        for _ in range(10):
            C = A @ B_t
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            C = A @ B_t
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return {"what": "cusparseSpMM, CSR, %s, through torch.sparse CSR @ dense (10 warm-ups, %d timed; baseline/base_cusparse/spmm.cu protocol)" % (str(B_t.dtype).replace("torch.", ""), steps),
                "ms_per_step": ms, "value": 2.0 * len(r) * B_t.shape[1] / (ms * 1e-3) / 1e9, "unit": "GFLOP/s", "note": "one matrix, A stays in L2 between launches where it fits"}
    except Exception as e:  # a yardstick must never break the line
        return {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU SpMM (spmm_reference_host, cuda_code/kernel_lib.hpp:859-881,
    compiled from /root/reference into oracle/_ref when it was built; else the oracle port) on all host threads."""
    if rank != 0:
        return
    from oracle import oracle_check as oc
    r, c, v = make_weight(1)
    N = args.n
    B = np.random.default_rng(7).uniform(-1, 1, (COLS, N)).astype(np.float32)
    indptr = oc.csr_from_coo(ROWS, r)
    cores = os.cpu_count() or 1
    if oc.ref_lib() is not None:
        kind = "reference"
        fn = lambda: oc.ref_spmm_f32(ROWS, N, COLS, indptr, c, v, B, threads=cores)
    else:
        kind = "port"
        fn = lambda: oc.oracle_spmm_f32(ROWS, N, indptr, c, v, B, mt=True)
    steps = max(1, min(args.steps, 200))
    for _ in range(max(1, min(args.warmup, 3))):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = (time.perf_counter() - t0) / steps
    gf = 2.0 * len(r) * N / dt / 1e9
    line = {"impl": "reference", "metric": "SpMM GFLOP/s (2*nnz*N/t)", "value": gf, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "pruned Llama-2-7B FFN weight 11008x4096, %.0f%% unstructured, N=%d (CPU: fp32)" % (100 * (1 - DENSITY), N), "nnz": int(len(r))},
            "cpu_baseline": {"value": gf, "unit": "GFLOP/s", "cores": cores, "kind": kind,
                             "sample": "the whole 11008x4096 matrix, %d passes, all host threads over row chunks" % steps},
            "e2e": {"value": gf, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def c4_block(args, rank, world, local, dist):
    """BASELINE.json configs[3]: synthetic power-law matrix (10 M rows, ~200 M nnz), N=32, fp32, cut into nnz-balanced
    contiguous row blocks with the reference's balanced-interval rule, one block per GPU; B replicated, C rows disjoint,
    no collective on the data path.  The matrix is fixed, so this is strong scaling.  Every rank calls this; rank 0 gets the dict."""
    import torch
    from generalsparse_b200 import api, shard, synth
    from oracle import oracle_check as oc

    N, rows = 32, args.rows
    hbm_peak, peak_src = peaks()
    t0 = time.perf_counter()
    lengths = synth.powerlaw_lengths(rows, seed=4, avg_deg=args.avg_deg, alpha=2.1, cap=min(1_000_000, rows // 5))
    cuts = shard.row_cuts(lengths, world)
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    r, c = synth.powerlaw_block(rows, rows, lengths, lo, hi, seed=4)
    v = ((((r + lo) * 2654435761 + c * 40503) % 2001).astype(np.float32) - 1000.0) * 1e-3  # a fixed function of (row, col) in [-1, 1]
    v[v == 0] = 0.5
    B = synth.synth_dense(rows, N, seed=4)
    gen_s = time.perf_counter() - t0
    my_rows, nnz = hi - lo, int(len(r))
    t0 = time.perf_counter()
    ms = api.MetaDataSet.from_coo(my_rows, rows, r, c, v)
    sess = api.Session(ms).run_family(args.family, N, 1, 1)
    host_pipeline_s = time.perf_counter() - t0
    eng = api.Engine(local)
    eng.upload_session(sess)
    if args.c4_kernel == "auto":
        c4_kernel = "b200_nnz" if args.family in ("thread_bit_map", "warp_segment") else "auto"
    else:
        c4_kernel = None if args.c4_kernel == "reference" else args.c4_kernel
    plan = eng.plan(sess, N, kernel=c4_kernel)
    plan.set_dense(B)
    dtf = None
    if args.family in ("thread_bit_map", "warp_total", "thread_total"):
        dtf = device_transform_block(eng, sess, args.family, my_rows, rows, r, c, v, N, 1, False, plan.desc.sub_matrix_id, host_pipeline_s, c4_kernel)
    dev = torch.device("cuda", local)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)
    # C of this rank lands in place in a full-size C (rows x N): what the optional all-gather replicates
    C_full = torch.zeros((rows, N), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    plan.set_output_device(C_full.data_ptr() + lo * N * 4)
    # parity on EVERY rank: the first rows of the block and the rows around the block's longest row (the hub row sits in one of the
    # blocks; the f64 oracle over all 200 M non-zeros would take minutes, rows are independent)
    plan.run(1)
    C = plan.output()
    per_row = np.bincount(r, minlength=my_rows)
    long_row = int(per_row.argmax())
    spans = [(0, min(my_rows, 20000)), (max(0, long_row - 8), min(my_rows, long_row + 8))]
    ok_local, checked, worst = True, 0, 0.0
    rowstart = np.concatenate([[0], np.cumsum(per_row)]).astype(np.int64)
    for a, b in spans:
        s0, s1 = int(rowstart[a]), int(rowstart[b])
        ip = (rowstart[a:b + 1] - s0).astype(np.int64)
        ref = oc.oracle_spmm_f64acc(b - a, N, ip, c[s0:s1], v[s0:s1], B)
        absrow = oc.oracle_spmm_absrow(b - a, N, ip, c[s0:s1], v[s0:s1], B)
        err = np.abs(C[a:b].astype(np.float64) - ref)
        ok_local = ok_local and bool((err <= 1e-5 * np.abs(ref) + 1e-6 * absrow).all())
        worst = max(worst, float((err / np.maximum(np.abs(ref), 1e-3 * absrow + 1e-30)).max()))
        checked += b - a
    ok = torch.tensor([1.0 if ok_local else 0.0, -float(per_row.max()), -worst], device=dev, dtype=torch.float64)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        e1.synchronize()
        barrier()
        return e0.elapsed_time(e1)

    for _ in range(max(3, min(args.warmup, 5))):
        plan.launch()
    steps = max(1, min(args.c4_steps, 200))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.15)
    tw0 = time.time()
    ms_spmm = timed(plan.launch, steps)
    timed(plan.launch, max(steps, int(600 / max(ms_spmm / steps, 1e-3))))  # ~0.6 s more of the same load for nvidia-smi to see (not reported)
    tw1 = time.time()
    clocks = sampler.stop(tw0, tw1) if rank == 0 else None
    # optional: C replicated on every GPU over NCCL (ragged blocks: one group of broadcasts, in place)
    ag = None
    if world > 1:
        ids = [api.Engine.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        eng.comm_init(world, rank, ids[0])
        cuts64 = np.ascontiguousarray(cuts, dtype=np.uint64)
        gather = lambda: eng.comm_allgather_rows(C_full.data_ptr(), cuts64, N * 4)
        both = lambda: (plan.launch(), gather())
        both()
        barrier()
        # every replica must equal what the owners computed: one checksum per rank, min and max over ranks must agree
        chk = C_full.double().sum().reshape(1)
        mn, mx = chk.clone(), chk.clone()
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        ms_both = timed(both, steps)
        ms_ag = timed(gather, steps)
        agt = torch.tensor([ms_both, ms_ag], device=dev, dtype=torch.float64)
        dist.all_reduce(agt, op=dist.ReduceOp.MAX)
        ag = {"api": "gs_comm_allgather_rows (NCCL, grouped ncclBroadcast of the ragged row blocks, in place in a rows x N C per GPU)",
              "ms_per_step_spmm_plus_allgather": float(agt[0].item()) / steps, "ms_allgather_alone": float(agt[1].item()) / steps,
              "bytes_gathered_per_gpu": int(rows * N * 4), "replicas_identical": bool(mn.item() == mx.item()) and bool(abs(mx.item()) > 0)}
    # whole-job SURVEY 8(d) bytes: every non-zero and offset once, B once and C once (NOT B once per GPU: replicating B is this
    # implementation's choice, not compulsory traffic), against world x the measured peak
    t = torch.tensor([ms_spmm, float(nnz), float(plan.algorithmic_bytes - (rows + my_rows) * N * 4)], device=dev, dtype=torch.float64)
    tmax = t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    kname, launches = plan.kernel_name, steps * plan.launch_count
    del plan, C_full
    yard = None
    if world == 1 and not args.no_yardstick:
        del eng
        torch.cuda.empty_cache()
        yard = cusparse_yardstick(my_rows, rows, r, c, v, torch.from_numpy(B).cuda(), steps=10)
    if rank != 0:
        return None
    ms_step = float(tmax[0].item()) / steps
    nnz_all = float(t[1].item())
    bytes_once = float(t[2].item()) + 2.0 * rows * N * 4
    tr = traffic_bytes(N, False, world, "c4")
    return {"config": "power-law matrix %d x %d, %.0f nnz, N=32, fp32, %s chain, ONE matrix cut into %d nnz-balanced row block(s) (strong scaling)" % (rows, rows, nnz_all, args.family, world),
            "value": 2.0 * nnz_all * N / (ms_step * 1e-3) / 1e9, "unit": "GFLOP/s", "ms_per_step": ms_step, "steps": steps, "scaling": "strong",
            "efficiency": "t(1) / (G * t(G)) with t = c4.ms_per_step of the N=1 and the N=G line", "kernel": kname, "gpu_launches": launches,
            "max_row_nnz": int(-ok[1].item()), "nnz_imbalance": float(tmax[1].item()) * world / nnz_all,
            "roofline": {"bound": "hbm", "algorithmic_bytes_whole_job": int(bytes_once), "achieved": bytes_once / (ms_step * 1e-3) / 1e9, "peak": hbm_peak * world, "unit": "GB/s",
                         "frac": bytes_once / (ms_step * 1e-3) / 1e9 / (hbm_peak * world),
                         "traffic": tr if world == 1 else None, "traffic_over_algorithmic": (tr / bytes_once if tr and world == 1 else None),
                         "note": "B and C counted once for the whole job; traffic = ncu dram bytes of one launch on 1 GPU (profiles/traffic.json, a committed capture)"},
            "parity": {"rows_checked_per_rank": checked, "includes_longest_row_of_every_block": True, "tolerance": "1e-5*|ref| + 1e-6*sum|a*b|", "worst_rel_err": float(-ok[2].item()),
                       "ok_on_every_rank": bool(ok[0].item() > 0.5)},
            "clocks": clocks, "allgather": ag, "yardstick_cusparse": yard, "generate_s": round(gen_s, 1), "host_pipeline_s": round(host_pipeline_s, 1), "device_transform": dtf}


def run_c4(args, rank, world, local):
    """--workload c4: the c4 block alone, as a line of its own (experiments; the driver-run line carries it as `c4`)"""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    blk = c4_block(args, rank, world, local, dist)
    if rank == 0:
        line = {"metric": "SpMM GFLOP/s (2*nnz*N/t)", "value": blk["value"] if blk["parity"]["ok_on_every_rank"] else None, "unit": "GFLOP/s", "n_gpus": world, "steps": blk["steps"],
                "warmup": 3, "ms_per_step": blk["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": blk["config"], "kernel": blk["kernel"]}, "roofline": blk["roofline"], "gpu_launches": blk["gpu_launches"], "clocks": blk["clocks"],
                "parity": blk["parity"], "c4": blk}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not blk["parity"]["ok_on_every_rank"]:
        sys.exit(1)


def main():
    global DENSITY, PATTERN
    args = parse()
    DENSITY, PATTERN = args.density, args.pattern
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    if args.workload == "c4":
        return run_c4(args, rank, world, local)

    import torch
    import torch.distributed as dist
    from generalsparse_b200 import api, synth

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N, half = args.n, True
    # --kernel auto: the LIBRARY picks (gs_plan_set_kernel(GS_KERNEL_AUTO): cost model over the chain, the matrix, N and the value type)
    hbm_peak, peak_src = peaks()

    # ---- the global matrix: `world` FFN-sized blocks stacked; this rank's nnz-balanced row block ------------------------
    blocks = [make_weight(1 + g) for g in range(world)]
    row_nnz = np.concatenate([np.bincount(b[0], minlength=ROWS) for b in blocks]).astype(np.uint64)
    from generalsparse_b200 import shard
    cuts = shard.row_cuts(row_nnz, world)  # the same rule tests/test_partition_gloo.py covers on CPU
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    g_rows = np.concatenate([b[0] + g * ROWS for g, b in enumerate(blocks)])
    sel = (g_rows >= lo) & (g_rows < hi)
    r = (g_rows[sel] - lo).astype(np.int64)
    c = np.concatenate([b[1] for b in blocks])[sel]
    v = np.concatenate([b[2] for b in blocks])[sel]
    my_rows = hi - lo
    nnz = int(len(r))
    del blocks, g_rows, sel

    # ---- the reference's operator chain (token_test.cc:1003-1092) with sparse coarsen 8 -------------------------------
    t_host0 = time.perf_counter()
    ms = api.MetaDataSet.from_coo(my_rows, COLS, r, c, v)
    sess = api.Session(ms).run_family("thread_total", N, 8, 1)
    host_pipeline_s = time.perf_counter() - t_host0
    eng = api.Engine(local)
    desc = sess.plan_desc()
    plans = []
    B_host = torch.empty((COLS, N), dtype=torch.float16).pin_memory()
    B_host.copy_(torch.from_numpy(np.random.default_rng(7).uniform(-1, 1, (COLS, N)).astype(np.float16)))
    C_host = torch.empty((desc.rows, N), dtype=torch.float16).pin_memory()
    copies = 1
    k = 0
    while k < copies:
        eng.upload_session(sess, half=half, as_id=k)
        p = eng.plan(sess, N, half=half, kernel=args.kernel, sub_matrix_id=k)
        p.set_dense(B_host.numpy())
        plans.append(p)
        if k == 0:  # enough device copies of A (in the layout the chosen kernel streams) to exceed L2 three times
            bytes_a = max(1, p.format_bytes - (COLS + int(desc.rows)) * N * 2)
            copies = args.copies or max(2, int(3 * 126e6 / bytes_a) + 1)
        k += 1
    del p
    kfam = "b200_sp24" if "sp24" in plans[0].kernel_name else "b200_tcgen05" if "k_b200_tile" in plans[0].kernel_name else "b200_row_smem" if "k_b200_" in plans[0].kernel_name else args.kernel
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))

    # ---- parity gate on the timed configuration, on EVERY rank (oracle = checker only) ----------------------------------
    # |C - C64| <= 1e-2 * |C64| + 1e-6 * sum|a*b| (+ 2^-25: half the spacing of fp16 subnormals)
    from oracle import oracle_check as oc
    plans[0].run(1)
    C = plans[0].output().astype(np.float64)
    indptr = np.concatenate([[0], np.cumsum(np.bincount(r, minlength=my_rows))]).astype(np.int64)
    Bf = B_host.numpy().astype(np.float32)
    ref = oc.oracle_spmm_f64acc(my_rows, N, indptr, c, v, Bf)
    absrow = oc.oracle_spmm_absrow(my_rows, N, indptr, c, v, Bf)
    err = np.abs(C - ref)
    par = torch.tensor([1.0 if bool((err <= 1e-2 * np.abs(ref) + 1e-6 * absrow + 2.0 ** -25).all()) else 0.0,
                        -float((err / np.maximum(np.abs(ref), 1e-3 * absrow + 1e-30)).max())], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(par, op=dist.ReduceOp.MIN)
    parity = {"max_rel_err_vs_f64": -float(par[1].item()), "tolerance": "1e-2*|ref| + 1e-6*sum|a*b| + 2^-25", "ok": bool(par[0].item() > 0.5),
              "ranks_checked": world}
    del C, ref, absrow, err
    dtf = device_transform_block(eng, sess, "thread_total", my_rows, COLS, r, c, v, N, 8, half, 0, host_pipeline_s, args.kernel)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            fn(i)
        e1.record(stream)
        e1.synchronize()
        barrier()
        ms_ = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms_, op=dist.ReduceOp.MAX)
        return float(ms_.item())

    launch = lambda i: plans[i % copies].launch()
    for i in range(max(3, args.warmup)):
        launch(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.15)
    t0 = time.time()
    ms_total = timed(launch, args.steps)
    # keep the device under the same load long enough for nvidia-smi to see it (not part of the reported time)
    sustain_steps = max(args.steps, int(0.5 / max(ms_total / args.steps * 1e-3, 1e-6)))
    ms_sustain = timed(launch, min(sustain_steps, 200000))
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if rank == 0 else None

    # ---- e2e: pinned host B -> device, kernel, C -> pinned host, every step ---------------------------------------------
    bp, cp = B_host.data_ptr(), C_host.data_ptr()
    e2e_fn = lambda i: plans[i % copies].spmm_host_ptr(bp, cp)
    for i in range(3):
        e2e_fn(i)
    e2e_steps = max(10, min(args.steps, 500))
    barrier()
    tw0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_fn(i)  # returns after the stream drained: the D2H result is on the host
    torch.cuda.synchronize()
    e2e_sync_s = time.perf_counter() - tw0
    # the same round trip through gs_spmm_host_async: every step still copies B in and C out, but the host does not wait
    # between steps; copies and launches of consecutive steps (different plans, one pinned C buffer each) overlap.
    # The timer stops after gs_synchronize(): every C is on the host.
    C_hosts = [torch.empty((desc.rows, N), dtype=torch.float16).pin_memory() for _ in range(copies)]
    cps = [t.data_ptr() for t in C_hosts]
    for i in range(copies):
        plans[i].spmm_host_async_ptr(bp, cps[i])
    eng.synchronize()
    e2e_ok = bool(all(torch.equal(t, C_host) for t in C_hosts))  # same A, same B: must equal the synchronous result
    barrier()
    tw0 = time.perf_counter()
    for i in range(e2e_steps):
        plans[i % copies].spmm_host_async_ptr(bp, cps[i % copies])
    e2e_enqueue_s = time.perf_counter() - tw0  # the host's share: all calls issued, the device may still be working
    eng.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - tw0, e2e_sync_s], device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s, e2e_sync_s = float(e2e_s[0].item()), float(e2e_s[1].item())

    # ---- dependent chain: step i+1 multiplies what step i produced (B := the first K rows of the previous C, a real data dependency
    # on one stream, no cross-launch overlap): what a pruned FFN layer at N tokens sees, beside the back-to-back figure above
    dep_steps = max(10, min(args.steps, 1000))

    def dep(i):
        plans[i % copies].set_dense_device(plans[(i - 1) % copies].output_device())
        plans[i % copies].launch()
    for i in range(1, 4):
        dep(i)
    ms_dep = timed(dep, dep_steps) / dep_steps
    # one launch alone: event pair around a single launch, the device idle before it (what ncu's gpu__time_duration sees, warm caches aside)
    iso = []
    for i in range(64):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        plans[i % copies].launch()
        e1.record(stream)
        e1.synchronize()
        iso.append(e0.elapsed_time(e1))
    ms_iso = float(np.median(iso))
    # the floor of that measurement: the same event pair around one launch of the same kernel family on a 256 x 256 matrix (what an idle device
    # charges for any launch: profiles/r2_sp24_timeline.md)
    tr, tc = synth.synth_matrix("uniform", 256, 256, seed=1, density=0.05)
    tsess = api.Session(api.MetaDataSet.from_coo(256, 256, tr, tc, None)).run_family("thread_total", N, 8, 1)
    eng.upload_session(tsess, half=half, as_id=901)
    tplan = eng.plan(tsess, N, half=half, kernel=kfam if kfam.startswith("b200") else args.kernel, sub_matrix_id=901)
    tplan.set_dense(np.ones((256, N), np.float16 if half else np.float32))
    for _ in range(5):
        tplan.launch()
    iso_floor = []
    for i in range(64):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        tplan.launch()
        e1.record(stream)
        e1.synchronize()
        iso_floor.append(e0.elapsed_time(e1))
    ms_iso_floor = float(np.median(iso_floor))
    del tplan, tsess

    nnz_all = torch.tensor([desc.origin_nnz], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(nnz_all)
    nnz_all = float(nnz_all.item())
    alg_bytes, fmt_bytes, kernel_name, launch_count = plans[0].algorithmic_bytes, plans[0].format_bytes, plans[0].kernel_name, plans[0].launch_count
    yard = None
    if world == 1 and not args.no_yardstick:
        yard = {"f16": cusparse_yardstick(my_rows, COLS, r, c, v, torch.from_numpy(B_host.numpy().copy()).cuda()),
                "f32": cusparse_yardstick(my_rows, COLS, r, c, v, torch.from_numpy(B_host.numpy().astype(np.float32)).cuda())}
    c4 = None
    if not args.no_c4:
        del plans, C_hosts, eng, launch, e2e_fn, dep
        torch.cuda.empty_cache()
        c4 = c4_block(args, rank, world, local, dist)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_step = ms_total / args.steps
    value = 2.0 * nnz_all * N / (ms_step * 1e-3) / 1e9
    achieved = alg_bytes / (ms_step * 1e-3) / 1e9
    args.kfam = kfam
    traffic = traffic_bytes(N, half, world, args)
    ok_all = parity["ok"] and (c4 is None or c4["parity"]["ok_on_every_rank"])
    line = {
        "metric": "SpMM GFLOP/s (2*nnz*N/t)", "value": value if ok_all else None, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 storage, f32 accumulate",
        "data": "synthetic",
        "config": {"workload": "pruned Llama-2-7B FFN weight 11008x4096 per GPU, %s, N=%d, HALF, thread_total chain (sort, S=8 pad) on the B200 %s kernel" % (
                       "2:4 structured" if args.pattern == "nm24" else "%.0f%% unstructured" % (100 * (1 - DENSITY)), N,
                       {"b200_row_smem": "row", "b200_tcgen05": "tile (tcgen05)", "b200_sp24": "2:4 (tcgen05.mma.sp)"}.get(kfam, kfam) + (" (picked by the library: --kernel auto)" if args.kernel == "auto" else "")),
                   "rows_per_gpu": int(desc.rows), "cols": COLS, "nnz_per_gpu": int(desc.origin_nnz), "stored_nnz_per_gpu": int(desc.stored_nnz),
                   "kernel": kernel_name, "l2": "steps rotate over %d device copies of A (%.0f MB > 126 MB L2)" % (copies, copies * bytes_a / 1e6),
                   "partition": "nnz-balanced row blocks (balanced_interval rule), B replicated, C rows disjoint, no collective",
                   "host_pipeline_s": round(host_pipeline_s, 3)},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                     "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback 6650",
                     "algorithmic_bytes_per_launch": int(alg_bytes), "algorithmic_bytes_rule": "SURVEY.md 8(d): stored nnz * (idx + val) + the chain's offset arrays + permutation + K*N + M*N values",
                     "format_bytes_per_launch": int(fmt_bytes), "format_over_algorithmic": fmt_bytes / alg_bytes,
                     "traffic_over_algorithmic": (traffic / alg_bytes if traffic else None),
                     "traffic_source": "profiles/traffic.json: one ncu --set full capture of this kernel, committed; not measured in this run",
                     "sustained_ms_per_step": ms_sustain / min(sustain_steps, 200000),
                     "dependent_chain_ms_per_step": ms_dep, "dependent_chain_frac": alg_bytes / (ms_dep * 1e-3) / 1e9 / hbm_peak,
                     "dependent_chain": "step i+1's B is the first K rows of step i's C: one stream; the launch starts under programmatic dependent launch and waits on the device (griddepcontrol.wait) before it requests B",
                     "isolated_launch_ms": ms_iso, "isolated_launch_frac": alg_bytes / (ms_iso * 1e-3) / 1e9 / hbm_peak,
                     "isolated_launch": "median of 64 single launches, device idle before each, CUDA events around the one launch",
                     "isolated_launch_floor_ms": ms_iso_floor,
                     "isolated_launch_floor": "the same event pair around one launch of the same kernel on a 256 x 256 matrix: what an idle device charges for any launch"},
        "e2e": {"value": 2.0 * nnz_all * N * e2e_steps / e2e_s / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": int(COLS * N * 2),
                "d2h_bytes_per_step": int(desc.rows * N * 2), "ms_per_step": e2e_s / e2e_steps * 1e3, "steps": e2e_steps,
                "api": "gs_spmm_host_async per step (pinned B in, pinned C out) + one gs_synchronize; results equal the blocking call: %s" % e2e_ok,
                "blocking_ms_per_step": e2e_sync_s / e2e_steps * 1e3, "host_enqueue_ms_per_step": e2e_enqueue_s / e2e_steps * 1e3},
        "gpu_launches": args.steps * launch_count, "clocks": clocks, "parity": parity, "c4": c4, "yardstick_cusparse": yard,
        "device_transform": dtf,
    }
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        if oc.ref_lib() is not None:
            kind, fn = "reference", (lambda: oc.ref_spmm_f32(my_rows, N, COLS, indptr, c, v, Bf, threads=cores))
        else:
            kind, fn = "port", (lambda: oc.oracle_spmm_f32(my_rows, N, indptr, c, v, Bf, mt=True))
        fn()
        best = 1e30
        for _ in range(5):
            tc = time.perf_counter()
            fn()
            best = min(best, time.perf_counter() - tc)
        line["cpu_baseline"] = {"value": 2.0 * nnz * N / best / 1e9, "unit": "GFLOP/s", "cores": cores, "kind": kind,
                                "sample": "the whole 11008x4096 matrix, best of 5 passes, fp32, all host threads over row chunks"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if not ok_all:
        sys.exit(1)


if __name__ == "__main__":
    main()
