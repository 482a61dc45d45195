"""C parity of the CUDA kernels with the oracle, through the C ABI, on a B200.

Gates (BASELINE.json north_star):
  * reference mode (all-ones A and B): EXACT equality with the oracle -- the reference's own check
    (code_generator.cc:631-638, cuda_code/kernel_lib.hpp:884-921: C[i][*] == nnz(row i));
  * random values: max |C - C64| <= tol * max(|C64|_inf-row scale) with tol 1e-5 (fp32) / 1e-2 (HALF,
    fp32 accumulate), plus the abs floor 1e-6 * sum|a*b| of the row.
"""
import glob
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD_DIR = os.path.join(os.path.dirname(__file__), "golden")
FAMILIES = ["thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment", "block_total", "block_bitmap"]


@pytest.fixture(scope="module")
def engine(built_lib):
    from generalsparse_b200 import api
    return api.Engine(0)


def _matrix(kind, rows, cols, seed, **kw):
    from generalsparse_b200 import synth
    r, c = synth.synth_matrix(kind, rows, cols, seed=seed, **kw)
    return r, c


def _check(C, ref64, absrow, tol, abs_floor=0.0):
    C = C.astype(np.float64)
    err = np.abs(C - ref64)
    # rel err on C with the row-L1 floor; abs_floor: half the spacing of fp16 subnormals when C is stored in fp16
    bound = tol * np.maximum(np.abs(ref64), 1e-30) + tol * absrow + abs_floor
    bad = err > bound
    assert not bad.any(), "max err %g at %s (ref %g, floor %g)" % (err.max(), np.argwhere(bad)[0], ref64[bad][0], absrow[bad][0])


def _run(engine, family, rows, cols, r, c, vals, B, N, S, C_, half, kernel=None):
    from generalsparse_b200 import api
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family(family, N, S, C_)
    engine.upload_session(s, half=half)
    plan = engine.plan(s, N, half=half, kernel=kernel)
    plan.set_dense(B)
    plan.run(1)
    return plan.output(), plan


CASES = [("tiny", "uniform", 64, 48, dict(density=0.085)), ("ragged", "ragged", 200, 150, dict(mean_len=9.0, empty_frac=0.15)),
         ("ffnlike", "uniform", 344, 512, dict(density=0.30)), ("powerlaw", "powerlaw", 3000, 3000, dict(avg_deg=20.0, alpha=2.1)),
         ("ig5like", "lognormal", 1500, 1300, dict(mean_len=37.4))]


@pytest.mark.parametrize("family", FAMILIES)
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("N,coarsen", [(8, 1), (32, 1), (16, 2), (5, 1)])
def test_reference_mode_exact(engine, family, case, N, coarsen):
    """all-ones A and B: C[i][*] == nnz(row i), exact, for every family on every matrix shape"""
    from generalsparse_b200 import api
    from oracle.oracle_check import csr_from_coo, oracle_spmm_f32
    name, kind, rows, cols, kw = case
    r, c = _matrix(kind, rows, cols, seed=CASES.index(case) + 1, **kw)
    B = np.ones((cols, N), dtype=np.float32)
    S = 4 if family == "thread_total" else 1
    try:
        C, plan = _run(engine, family, rows, cols, r, c, None, B, N, S, coarsen, half=False)
    except api.GsError as e:
        if e.status == 5:  # PADDING_RATE_UP_BOUND: the reference aborts on the same input (golden SUMMARY.json)
            pytest.skip("padding rate bound, as in the reference")
        raise
    ref = oracle_spmm_f32(rows, N, csr_from_coo(rows, r), c, np.ones(len(r), np.float32), B)
    assert np.array_equal(C, ref)
    assert np.array_equal(C[:, 0], np.bincount(r, minlength=rows).astype(np.float32))


@pytest.mark.parametrize("family", FAMILIES)
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
def test_random_values_tolerance(engine, family, half):
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    rows, cols, N = 344, 512, 16
    r, c = _matrix("uniform", rows, cols, seed=4, density=0.3)
    vals = synth.synth_values(len(r), seed=4)
    B = synth.synth_dense(cols, N, seed=4)
    if half:  # the oracle sees exactly the fp16-rounded inputs the device stores
        vals = vals.astype(np.float16).astype(np.float32)
        B = B.astype(np.float16).astype(np.float32)
    S = 4 if family == "thread_total" else 1
    C, plan = _run(engine, family, rows, cols, r, c, vals, B, N, S, 1, half=half)
    indptr = csr_from_coo(rows, r)
    ref = oracle_spmm_f64acc(rows, N, indptr, c, vals, B)
    absrow = oracle_spmm_absrow(rows, N, indptr, c, vals, B)
    if half and family in ("thread_total", "warp_total", "block_total"):
        # fp32 accumulation, one final rounding to fp16 (2^-11 relative): SURVEY.md 8(d)'s gate, floor 1e-6 * |row|_1 + half the
        # spacing of fp16 subnormals
        _check(C, ref, absrow * 1e-4, 1e-2, abs_floor=2.0 ** -25)
    elif half:
        # the bit-map / segment families add their partial sums into an fp16 C with atomicAdd(half), as the reference's HALF tokens
        # do (total_BMT...token.cc:510-513): every partial is rounded to fp16 on the way, so the error scales with the row's
        # L1 norm, not with |C| -- the loose floor stays for these four families only
        _check(C, ref, absrow, 1e-2)
    else:
        _check(C, ref, absrow * 1e-1, 1e-5)


@pytest.mark.parametrize("N", [8, 16, 32])
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
@pytest.mark.parametrize("chain", ["thread_total", "warp_total", "block_total"])
def test_b200_row_kernel(engine, N, half, chain):
    """the B200 row kernel on each row-direction chain: exact in reference mode, in tolerance on random values"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    rows, cols = 1000, 4096 if N <= 16 else 1024
    r, c = _matrix("uniform", rows, cols, seed=21, density=0.1)
    r2, c2 = _matrix("ragged", 300, cols, seed=22, mean_len=20.0, empty_frac=0.2)
    for (rr, cc, nrows) in ((r, c, rows), (r2, c2, 300)):
        indptr = csr_from_coo(nrows, rr)
        S = 8 if chain == "thread_total" else 1
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, None, B1, N, S, 1, half=half, kernel="b200_row_smem")
        assert "k_b200_" in plan.kernel_name
        ref = oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)
        assert np.array_equal(C.astype(np.float32), ref), plan.kernel_name
        vals = synth.synth_values(len(rr), seed=5)
        B = synth.synth_dense(cols, N, seed=5)
        if half:
            vals = vals.astype(np.float16).astype(np.float32)
            B = B.astype(np.float16).astype(np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, vals, B, N, S, 1, half=half, kernel="b200_row_smem")
        ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
        absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
        _check(C, ref, absrow * (1e-4 if half else 1e-1), 1e-2 if half else 1e-5, abs_floor=2.0 ** -25 if half else 0.0)


def test_host_roundtrip_and_execute(engine):
    """gs_spmm_host (H2D B, launch, D2H C) and gs_execute (execute_binary's timing + gflops formula)"""
    from generalsparse_b200 import api
    rows, cols, N = 344, 512, 8
    r, c = _matrix("uniform", rows, cols, seed=4, density=0.3)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    engine.upload_session(s)
    plan = engine.plan(s, N, kernel="b200_row_smem")
    B = np.ones((cols, N), dtype=np.float32)
    C = plan.spmm_host(B)
    assert np.array_equal(C[:, 0], np.bincount(r, minlength=rows).astype(np.float32))
    ms_total, gflops = plan.execute(20)
    assert ms_total > 0 and abs(gflops - 2.0 * len(r) * N * 20 / (ms_total * 1e-3) / 1e9) / gflops < 1e-3
    assert plan.algorithmic_bytes > len(r) * 6


def test_async_roundtrip_equals_blocking(engine):
    """gs_spmm_host_async over several plans (copy-in / launch / copy-out streams chained by events) == gs_spmm_host;
    back-to-back row-kernel launches also exercise programmatic dependent launch with different B per step"""
    import torch
    from generalsparse_b200 import api, synth
    rows, cols, N = 1000, 4096, 8
    r, c = _matrix("uniform", rows, cols, seed=31, density=0.1)
    vals = synth.synth_values(len(r), seed=31).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    plans = []
    for k in range(3):
        engine.upload_session(s, half=True, as_id=20 + k)
        plans.append(engine.plan(s, N, half=True, kernel="b200_row_smem", sub_matrix_id=20 + k))
    Bs = [torch.from_numpy(synth.synth_dense(cols, N, seed=40 + i).astype(np.float16)).pin_memory() for i in range(6)]
    want = [plans[i % 3].spmm_host(Bs[i].numpy()) for i in range(6)]
    outs = [torch.empty((rows, N), dtype=torch.float16).pin_memory() for _ in range(6)]
    for rep in range(3):  # queue depth 6, three times over
        for i in range(6):
            plans[i % 3].spmm_host_async_ptr(Bs[i].data_ptr(), outs[i].data_ptr())
    engine.synchronize()
    for i in range(6):
        assert np.array_equal(outs[i].numpy(), want[i]), i


@pytest.mark.parametrize("kernel,density", [("b200_sp24", 0.3), ("b200_tcgen05", 0.1)])
def test_async_roundtrip_rotates_over_launch_streams(engine, kernel, density):
    """gs_spmm_host_async puts consecutive steps on different launch streams: kernels that hand split panels over between
    their CTAs (every CTA of a launch resident) run side by side, a plan's own launches stay in order (5 plans over 4
    streams: every plan meets every stream), and a launch on gs_stream() right after the calls -- no gs_synchronize in
    between -- is ordered behind them"""
    import torch
    from generalsparse_b200 import api, synth
    rows, cols, N = 3000, 4096, 16
    r, c = _matrix("uniform", rows, cols, seed=33, density=density)
    vals = synth.synth_values(len(r), seed=33).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    plans = []
    for k in range(5):
        engine.upload_session(s, half=True, as_id=30 + k)
        plans.append(engine.plan(s, N, half=True, kernel=kernel, sub_matrix_id=30 + k))
    Bs = [torch.from_numpy(synth.synth_dense(cols, N, seed=50 + i).astype(np.float16)).pin_memory() for i in range(10)]
    want = [plans[i % 5].spmm_host(Bs[i].numpy()) for i in range(10)]
    outs = [torch.empty((rows, N), dtype=torch.float16).pin_memory() for _ in range(10)]
    for rep in range(20):
        for i in range(10):
            plans[i % 5].spmm_host_async_ptr(Bs[i].data_ptr(), outs[i].data_ptr())
    # the last async step of plan 4 used Bs[9]; a plain launch of the same plan must see that B and come after it
    plans[4].launch()
    last = plans[4].output()
    engine.synchronize()
    for i in range(10):
        assert np.array_equal(outs[i].numpy(), want[i]), i
    assert np.array_equal(last, want[9])


def test_errors_are_statuses_not_aborts(engine):
    from generalsparse_b200 import api
    rows, cols = 64, 48
    r, c = _matrix("uniform", rows, cols, seed=1, density=0.085)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c)
    s = api.Session(ms)
    with pytest.raises(api.GsError) as ei:  # pad-to-64 on ~4 nnz rows: the reference aborts, we return a status
        s.run_family("warp_bitmap", 8, 1, 1)
    assert ei.value.status == 5


@pytest.mark.parametrize("N", [16, 32, 64, 128, 256])
@pytest.mark.parametrize("chain", ["thread_total", "warp_total"])
def test_b200_tile_kernel(engine, N, chain):
    """decompress-to-shared-memory + tcgen05 path (HALF): exact in reference mode, in tolerance on random values;
    1000 x 4096 splits its 8 row panels over all CTAs (stream-K hand-off), 300 x 1000 has a ragged K tail and
    empty rows, 130 x 70 is smaller than two tiles"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    mats = [("uniform", 1000, 4096, dict(density=0.1)), ("ragged", 300, 1000, dict(mean_len=20.0, empty_frac=0.2)), ("uniform", 130, 70, dict(density=0.3))]
    for i, (kind, nrows, cols, kw) in enumerate(mats):
        rr, cc = _matrix(kind, nrows, cols, seed=51 + i, **kw)
        indptr = csr_from_coo(nrows, rr)
        S = 8 if chain == "thread_total" else 1
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, None, B1, N, S, 1, half=True, kernel="b200_tcgen05")
        assert "k_b200_tile" in plan.kernel_name
        ref = oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)
        assert np.array_equal(C.astype(np.float32), ref), plan.kernel_name
        vals = synth.synth_values(len(rr), seed=5).astype(np.float16).astype(np.float32)
        B = synth.synth_dense(cols, N, seed=5).astype(np.float16).astype(np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, vals, B, N, S, 1, half=True, kernel="b200_tcgen05")
        ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
        absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
        _check(C, ref, absrow * 1e-4, 1e-2, abs_floor=2.0 ** -25)
        # repeated launches (programmatic dependent launch, workspace and flag reuse) give the same bits
        plan.run(5)
        assert np.array_equal(plan.output(), C)


def test_b200_tile_overlapped_launches_keep_order(engine):
    """back-to-back launches of one plan overlap on the SMs (programmatic dependent launch, two CTAs per SM): C must end
    up as the LAST launch's product although the dense matrix alternates between launches"""
    import torch
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    rows, cols, N = 3000, 4096, 32
    r, c = _matrix("uniform", rows, cols, seed=61, density=0.2)
    vals = synth.synth_values(len(r), seed=61).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    engine.upload_session(s, half=True)
    plan = engine.plan(s, N, half=True, kernel="b200_tcgen05")
    Bs = [synth.synth_dense(cols, N, seed=70 + i).astype(np.float16) for i in range(2)]
    dev = [torch.from_numpy(b).cuda() for b in Bs]
    torch.cuda.synchronize()
    indptr = csr_from_coo(rows, r)
    refs = [oracle_spmm_f64acc(rows, N, indptr, c, vals, b.astype(np.float32)) for b in Bs]
    absrow = [oracle_spmm_absrow(rows, N, indptr, c, vals, b.astype(np.float32)) for b in Bs]
    for n_launch in (1, 2, 7, 40, 41):
        for i in range(n_launch):
            plan.set_dense_device(dev[i % 2].data_ptr())
            plan.launch()
        last = (n_launch - 1) % 2
        C = plan.output()
        _check(C, refs[last], absrow[last] * 1e-4, 1e-2, abs_floor=2.0 ** -25)


@pytest.mark.parametrize("N", [32, 64, 128])
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
@pytest.mark.parametrize("chain", ["thread_bit_map", "warp_segment"])
def test_b200_nnz_kernel(engine, N, half, chain):
    """the B200 nnz kernel (a warp per run of 32-nnz BMTs) on the nnz-direction bit-map chains: exact in reference mode,
    in tolerance on random values; power-law rows (rows far longer than a run, atomics), ragged rows with empty rows
    (segment_empty_flag / segment_empty_row_indices), a matrix smaller than one run"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    mats = [("powerlaw", 3000, 3000, dict(avg_deg=20.0, alpha=2.1)), ("ragged", 700, 450, dict(mean_len=9.0, empty_frac=0.3)),
            ("uniform", 40, 30, dict(density=0.1)), ("uniform", 344, 512, dict(density=0.3))]
    for i, (kind, nrows, cols, kw) in enumerate(mats):
        rr, cc = _matrix(kind, nrows, cols, seed=81 + i, **kw)
        indptr = csr_from_coo(nrows, rr)
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, None, B1, N, 1, 1, half=half, kernel="b200_nnz")
        assert "k_b200_nnz" in plan.kernel_name
        ref = oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)
        if not half or ref.max() <= 2048:  # fp16 C holds integers exactly up to 2048
            assert np.array_equal(C.astype(np.float32), ref), plan.kernel_name
        vals = synth.synth_values(len(rr), seed=5)
        B = synth.synth_dense(cols, N, seed=5)
        if half:
            vals = vals.astype(np.float16).astype(np.float32)
            B = B.astype(np.float16).astype(np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, vals, B, N, 1, 1, half=half, kernel="b200_nnz")
        ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
        absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
        floor = 2.0 ** -25 if half else 0.0  # half the spacing of fp16 subnormals when C is stored in fp16
        # HALF: rows that cross a run edge are completed with red.global.add on the fp16 C (each partial rounded to fp16), hence the
        # row-L1 floor at the tolerance itself; fp32 keeps SURVEY.md 8(d)'s 1e-6 * |row|_1
        _check(C, ref, absrow * (1.0 if half else 1e-1), 1e-2 if half else 1e-5, abs_floor=floor)
        plan.run(3)  # C is zeroed per launch: repeated launches do not accumulate
        _check(plan.output(), ref, absrow * (1.0 if half else 1e-1), 1e-2 if half else 1e-5, abs_floor=floor)


def test_b200_nnz_rejects_narrow_dense(engine):
    from generalsparse_b200 import api
    r, c = _matrix("uniform", 64, 48, seed=1, density=0.085)
    s = api.Session(api.MetaDataSet.from_coo(64, 48, r, c)).run_family("thread_bit_map", 8, 1, 1)
    engine.upload_session(s)
    with pytest.raises(api.GsError):
        engine.plan(s, 8, kernel="b200_nnz")


@pytest.mark.parametrize("N", [8, 16, 32, 64, 128, 256])
@pytest.mark.parametrize("chain", ["thread_total", "warp_total"])
def test_b200_sp24_kernel(engine, N, chain):
    """2:4 sparse tensor-core path (HALF): exact in reference mode, in tolerance on random values.  Exactly 2:4-pruned
    matrices have no remainder (one launch); unstructured ones send what the pattern cannot hold through remainder
    warps of the same kernel (N <= 64: tile-local blocks that travel with the tile images) or through the gather pass (two launches).  1000 x 4096 splits its 8 row panels over all CTAs (stream-K hand-off), 300 x 1000 has a
    ragged K tail and empty rows, 130 x 72 is smaller than two tiles"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    mats = [("nm24", 1000, 4096, {}, 1), ("nm24", 130, 72, {}, 1), ("uniform", 1000, 4096, dict(density=0.3), 2),
            ("ragged", 300, 1000, dict(mean_len=20.0, empty_frac=0.2), None)]
    for i, (kind, nrows, cols, kw, launches) in enumerate(mats):
        rr, cc = _matrix(kind, nrows, cols, seed=81 + i, **kw)
        indptr = csr_from_coo(nrows, rr)
        S = 8 if chain == "thread_total" else 1
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, None, B1, N, S, 1, half=True, kernel="b200_sp24")
        assert "k_b200_sp24" in plan.kernel_name
        if launches is not None:  # up to N=64 the remainder is multiplied by warps of the 2:4 kernel itself: one launch
            assert plan.launch_count == (1 if N <= 64 else launches), plan.kernel_name
            assert ("tile-local" in plan.kernel_name) == (launches == 2 and N <= 64), plan.kernel_name
        ref = oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)
        assert np.array_equal(C.astype(np.float32), ref), plan.kernel_name
        vals = synth.synth_values(len(rr), seed=5).astype(np.float16).astype(np.float32)
        B = synth.synth_dense(cols, N, seed=5).astype(np.float16).astype(np.float32)
        C, plan = _run(engine, chain, nrows, cols, rr, cc, vals, B, N, S, 1, half=True, kernel="b200_sp24")
        ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
        absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
        _check(C, ref, absrow * 1e-4, 1e-2, abs_floor=2.0 ** -25)
        # repeated launches (programmatic dependent launch, workspace, flag and remainder-buffer reuse) give the same bits
        plan.run(5)
        assert np.array_equal(plan.output(), C)


@pytest.mark.parametrize("N,panels", [(32, 2), (64, 1), (128, 1), (16, 2), (8, 2)])
def test_b200_sp24_panels_per_cta(engine, monkeypatch, N, panels):
    """the other unit size than the default for this N (GS_SP24_PANELS): one or two row panels per CTA and k step;
    9 panels is an odd count (the image is padded to whole units), with and without a remainder"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    monkeypatch.setenv("GS_SP24_PANELS", str(panels))
    for i, (kind, kw) in enumerate([("nm24", {}), ("uniform", dict(density=0.3))]):
        nrows, cols = 1100, 2048
        rr, cc = _matrix(kind, nrows, cols, seed=97 + i, **kw)
        indptr = csr_from_coo(nrows, rr)
        vals = synth.synth_values(len(rr), seed=6).astype(np.float16).astype(np.float32)
        B = synth.synth_dense(cols, N, seed=6).astype(np.float16).astype(np.float32)
        C, plan = _run(engine, "thread_total", nrows, cols, rr, cc, vals, B, N, 8, 1, half=True, kernel="b200_sp24")
        assert ("%d panel" % panels) in plan.kernel_name, plan.kernel_name
        ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
        absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
        _check(C, ref, absrow * 1e-4, 1e-2, abs_floor=2.0 ** -25)
        plan.run(7)
        assert np.array_equal(plan.output(), C)


@pytest.mark.parametrize("N,mode", [(8, 0), (16, 0), (32, 0), (8, 1), (16, 1), (32, 1)])
def test_b200_sp24_remainder_as_separate_pass(engine, monkeypatch, N, mode):
    """GS_SP24_FUSED_REM=0: the gather pass in front of the 2:4 launch also at the widths that fuse it by default;
    =1: gathers by eight warps of the CTA that owns the panel (N <= 16), not the tile-local blocks"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    monkeypatch.setenv("GS_SP24_FUSED_REM", str(mode))
    nrows, cols = 1000, 4096
    rr, cc = _matrix("uniform", nrows, cols, seed=99, density=0.3)
    vals = synth.synth_values(len(rr), seed=7).astype(np.float16).astype(np.float32)
    B = synth.synth_dense(cols, N, seed=7).astype(np.float16).astype(np.float32)
    C, plan = _run(engine, "thread_total", nrows, cols, rr, cc, vals, B, N, 8, 1, half=True, kernel="b200_sp24")
    assert plan.launch_count == (1 if mode == 1 and N <= 16 else 2) and "tile-local" not in plan.kernel_name, plan.kernel_name
    assert ("owner CTA" in plan.kernel_name) == (mode == 1 and N <= 16), plan.kernel_name
    indptr = csr_from_coo(nrows, rr)
    _check(C, oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B), oracle_spmm_absrow(nrows, N, indptr, cc, vals, B) * 1e-4, 1e-2, abs_floor=2.0 ** -25)
    plan.run(6)
    assert np.array_equal(plan.output(), C)


@pytest.mark.parametrize("N", [8, 16, 32, 64])
def test_b200_sp24_tile_local_remainder(engine, N):
    """the remainder cut along the tiles: dense (a stage of the ring must hold the largest block: 42 % keeps two CTAs per
    SM with three stages, 90 % does not fit and falls back), ragged K tail and row tail, a split panel per CTA (3000 x 4096: 24 panels over
    148 CTAs, every CTA leaves or collects partials that carry a share of the remainder), and a matrix whose one dense
    corner would make every unit as large as the largest (falls back as well); reference-mode ones exact"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    rng = np.random.default_rng(5)
    corner = rng.random((2000, 2048)) < 0.05
    corner[:128, :128] = True
    cr, ccol = np.nonzero(corner)
    mats = [("uniform", 3000, 4096, dict(density=0.3), True), ("uniform", 700, 1000, dict(density=0.42), True), ("uniform", 300, 520, dict(density=0.9), False),
            ("uniform", 130, 72, dict(density=0.4), True), (None, 2000, 2048, None, False)]
    for i, (kind, nrows, cols, kw, tiled) in enumerate(mats):
        rr, cc = _matrix(kind, nrows, cols, seed=31 + i, **kw) if kind else (cr.astype(np.int64), ccol.astype(np.int64))
        indptr = csr_from_coo(nrows, rr)
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _run(engine, "thread_total", nrows, cols, rr, cc, None, B1, N, 8, 1, half=True, kernel="b200_sp24")
        assert ("tile-local" in plan.kernel_name) == tiled, plan.kernel_name
        if tiled:
            assert plan.launch_count == 1 and "2 CTA/SM" in plan.kernel_name, plan.kernel_name
            assert ("%d stages" % (2 if N == 64 else 3 if N == 32 or i == 1 else 4)) in plan.kernel_name, plan.kernel_name
        assert np.array_equal(C.astype(np.float32), oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)), plan.kernel_name
        vals = synth.synth_values(len(rr), seed=8).astype(np.float16).astype(np.float32)
        B = synth.synth_dense(cols, N, seed=8).astype(np.float16).astype(np.float32)
        C, plan = _run(engine, "thread_total", nrows, cols, rr, cc, vals, B, N, 8, 1, half=True, kernel="b200_sp24")
        _check(C, oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B), oracle_spmm_absrow(nrows, N, indptr, cc, vals, B) * 1e-4, 1e-2, abs_floor=2.0 ** -25)
        plan.run(9)
        assert np.array_equal(plan.output(), C)


@pytest.mark.parametrize("N", [8, 16, 32, 64])
def test_b200_sp24_skips_empty_tiles(engine, monkeypatch, N):
    """tiles that hold nothing leave the stream (unit -> (panel, k tile) table): a block-sparse matrix (30 % dense blocks, the
    rest empty, some row panels entirely empty), a full band, and the same matrices with GS_SP24_SKIP_EMPTY=0;
    reference-mode ones exact"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    rng = np.random.default_rng(17)
    nrows, cols = 3000, 4096
    mask = rng.random((nrows, cols)) < 0.3
    blocks = rng.random((-(-nrows // 128), cols // 128)) < 0.25
    blocks[5] = False  # an empty row panel keeps one (empty) tile
    blocks[7, :] = False
    blocks[7, 31] = True  # a panel whose only tile is the last k tile
    mask &= np.kron(blocks, np.ones((128, 128), dtype=bool))[:nrows]
    br, bc = np.nonzero(mask)
    mats = [(br.astype(np.int64), bc.astype(np.int64), nrows, cols)]
    r2, c2 = _matrix("banded", 2000, 2048, seed=3, half_bandwidth=64)
    mats.append((r2, c2, 2000, 2048))
    for rr, cc, nr, nc in mats:
        indptr = csr_from_coo(nr, rr)
        vals = synth.synth_values(len(rr), seed=9).astype(np.float16).astype(np.float32)
        B = synth.synth_dense(nc, N, seed=9).astype(np.float16).astype(np.float32)
        B1 = np.ones((nc, N), dtype=np.float32)
        C1, plan = _run(engine, "thread_total", nr, nc, rr, cc, None, B1, N, 8, 1, half=True, kernel="b200_sp24")
        assert " tiles" in plan.kernel_name, plan.kernel_name
        assert np.array_equal(C1.astype(np.float32), oracle_spmm_f32(nr, N, indptr, cc, np.ones(len(rr), np.float32), B1)), plan.kernel_name
        C, plan = _run(engine, "thread_total", nr, nc, rr, cc, vals, B, N, 8, 1, half=True, kernel="b200_sp24")
        _check(C, oracle_spmm_f64acc(nr, N, indptr, cc, vals, B), oracle_spmm_absrow(nr, N, indptr, cc, vals, B) * 1e-4, 1e-2, abs_floor=2.0 ** -25)
        plan.run(9)
        assert np.array_equal(plan.output(), C)
        monkeypatch.setenv("GS_SP24_SKIP_EMPTY", "0")
        Cfull, plan = _run(engine, "thread_total", nr, nc, rr, cc, vals, B, N, 8, 1, half=True, kernel="b200_sp24")
        monkeypatch.delenv("GS_SP24_SKIP_EMPTY")
        assert " tiles" not in plan.kernel_name, plan.kernel_name  # another split of the panels over the CTAs: other fp32 summation order
        _check(Cfull, oracle_spmm_f64acc(nr, N, indptr, cc, vals, B), oracle_spmm_absrow(nr, N, indptr, cc, vals, B) * 1e-4, 1e-2, abs_floor=2.0 ** -25)


@pytest.mark.parametrize("kind,kw,N", [("nm24", {}, 32), ("uniform", dict(density=0.3), 32), ("uniform", dict(density=0.3), 16)])
def test_b200_sp24_overlapped_launches_keep_order(engine, kind, kw, N):
    """back-to-back launches of one plan overlap on the SMs: C must end up as the LAST launch's product although the
    dense matrix alternates between launches (with a remainder: two product buffers alternate as well)"""
    import torch
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    rows, cols = 3000, 4096
    r, c = _matrix(kind, rows, cols, seed=91, **kw)
    vals = synth.synth_values(len(r), seed=61).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    engine.upload_session(s, half=True)
    plan = engine.plan(s, N, half=True, kernel="b200_sp24")
    Bs = [synth.synth_dense(cols, N, seed=70 + i).astype(np.float16) for i in range(2)]
    dev = [torch.from_numpy(b).cuda() for b in Bs]
    torch.cuda.synchronize()
    indptr = csr_from_coo(rows, r)
    refs = [oracle_spmm_f64acc(rows, N, indptr, c, vals, b.astype(np.float32)) for b in Bs]
    absrow = [oracle_spmm_absrow(rows, N, indptr, c, vals, b.astype(np.float32)) for b in Bs]
    for n_launch in (1, 2, 7, 40, 41):
        for i in range(n_launch):
            plan.set_dense_device(dev[i % 2].data_ptr())
            plan.launch()
        last = (n_launch - 1) % 2
        C = plan.output()
        _check(C, refs[last], absrow[last] * 1e-4, 1e-2, abs_floor=2.0 ** -25)


def test_b200_sp24_metadata_lands_in_tensor_memory_as_laid_out(engine, monkeypatch):
    """the 4 metadata columns tcgen05.cp leaves in tensor memory for the first tile are the 16-byte rows of the image"""
    from generalsparse_b200 import api
    monkeypatch.setenv("GS_SP24_PROBE", "1")
    rows, cols, N = 128, 128, 16
    r, c = _matrix("nm24", rows, cols, seed=95)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c)
    s = api.Session(ms).run_family("warp_total", N, 1, 1)
    engine.upload_session(s, half=True)
    plan = engine.plan(s, N, half=True, kernel="b200_sp24")
    plan.set_dense(np.ones((cols, N), dtype=np.float32))
    plan.run(1)
    got = plan.sp24_probe()
    # expected: word (L, j): low half = groups k&15 of row (L&7) + 16*(L>>4), k in [32 j + 16*((L>>3)&1), +16); high half = row + 8
    keep = np.zeros((rows, cols), dtype=bool)
    keep[r, c] = True
    want = np.zeros((128, 4), dtype=np.uint32)
    for L in range(128):
        for j in range(4):
            w = 0
            for half in range(2):
                row = (L & 7) + 16 * (L >> 4) + 8 * half
                k0 = 32 * j + 16 * ((L >> 3) & 1)
                for q in range(4):
                    i0, i1 = np.nonzero(keep[row, k0 + 4 * q:k0 + 4 * q + 4])[0]
                    w |= (int(i0) | (int(i1) << 2)) << (16 * half + 4 * q)
            want[L, j] = w
    assert np.array_equal(got, want)


@pytest.mark.parametrize("N", [8, 16, 32])
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
def test_b200_row_short_rows(engine, monkeypatch, N, half):
    """(opt-in variant, GS_ROW_GROUP) short rows with a dense panel too large for shared memory (BASELINE's C1 shape in small): a group of 4 / 8 / 16 lanes per row, several rows
    in flight per warp; sorted (thread_total chain) and unsorted (warp_total chain) rows, empty rows, one long row; exact on ones, 8(d) gate on values"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    monkeypatch.setenv("GS_ROW_GROUP", "8")
    cols = 60000
    mats = [("lognormal", 3000, dict(mean_len=37.4)), ("ragged", 2500, dict(mean_len=9.0, empty_frac=0.3)), ("lognormal", 900, dict(mean_len=150.0))]
    for i, (kind, nrows, kw) in enumerate(mats):
        rr, cc = _matrix(kind, nrows, cols, seed=41 + i, **kw)
        indptr = csr_from_coo(nrows, rr)
        for chain, S in (("thread_total", 4), ("warp_total", 1)):
            B1 = np.ones((cols, N), dtype=np.float32)
            C, plan = _run(engine, chain, nrows, cols, rr, cc, None, B1, N, S, 1, half=half, kernel="b200_row_smem")
            assert "k_b200_row_short" in plan.kernel_name, plan.kernel_name
            ref = oracle_spmm_f32(nrows, N, indptr, cc, np.ones(len(rr), np.float32), B1)
            assert np.array_equal(C.astype(np.float32), ref), plan.kernel_name
            vals = synth.synth_values(len(rr), seed=5)
            B = synth.synth_dense(cols, N, seed=5)
            if half:
                vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
            C, plan = _run(engine, chain, nrows, cols, rr, cc, vals, B, N, S, 1, half=half, kernel="b200_row_smem")
            ref = oracle_spmm_f64acc(nrows, N, indptr, cc, vals, B)
            absrow = oracle_spmm_absrow(nrows, N, indptr, cc, vals, B)
            _check(C, ref, absrow * (1e-4 if half else 1e-1), 1e-2 if half else 1e-5, abs_floor=2.0 ** -25 if half else 0.0)
