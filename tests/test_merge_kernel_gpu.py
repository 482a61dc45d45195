"""The B200 range kernel (GS_FAMILY_B200_MERGE): the device consumer of merge-path, nnz-balanced, column-direction and multi-row-block
chains (round 1 produced their host arrays bit-exactly but nothing on the device read them).  Parity with the oracle through the C ABI:
all-ones exact, random values within SURVEY.md 8(d)'s gate.  Rows shared between segments are completed by tickets (fp32 partials, summed in
stream order by the last segment to arrive) unless a row has more than 64 sharers, where the plan keeps red.add into a zeroed C (each partial
rounded to fp16 when C is HALF, hence the row-L1 floor at the tolerance for HALF in _check)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(built_lib):
    from generalsparse_b200 import api
    return api.Engine(0)


def _mats():
    from generalsparse_b200 import synth
    return [("powerlaw", 3000, 3000, synth.synth_matrix("powerlaw", 3000, 3000, seed=5, avg_deg=20.0, alpha=2.1)),
            ("ragged", 700, 450, synth.synth_matrix("ragged", 700, 450, seed=2, mean_len=9.0, empty_frac=0.3)),
            ("tiny", 64, 48, synth.synth_matrix("uniform", 64, 48, seed=1, density=0.085))]


CHAINS = {
    "merge_path_thread": lambda N: [("merge_path_thread_operator", 33), ("thread_total_reduce_operator", 0, 1, 1), ("grid_block_operator", 64, min(N, 32), 4, 1)],
    "merge_path_warp": lambda N: [("merge_path_warp_operator", 300), ("warp_total_reduce_operator", 1), ("grid_block_operator", 64, 32, 4, 1)],
    "merge_path_tblock": lambda N: [("merge_path_tblock_operator", 4096), ("tblock_total_reduce_operator", 1), ("grid_block_operator", 64, 32, 4, 1)],
    "sort_merge_path_thread": lambda N: [("sort_operator",), ("merge_path_thread_operator", 7), ("thread_total_reduce_operator", 0, 1, 1), ("grid_block_operator", 64, min(N, 32), 4, 1)],
    "balanced_thread": lambda N: [("balanced_interval_row_direction_thread_blocking_operator", 40, 0, 0), ("thread_total_reduce_operator", 0, 1, 1), ("grid_block_operator", 64, min(N, 32), 4, 1)],
    "sort_balanced_warp": lambda N: [("sort_operator",), ("balanced_interval_row_direction_warp_blocking_operator", 500, 0, 0), ("warp_total_reduce_operator", 1), ("grid_block_operator", 64, 32, 4, 1)],
    "rowblock4_thread": lambda N: [("fixed_interval_row_direction_thread_blocking_operator", 4, 0, 0, 0, 0, 0, 1), ("thread_total_reduce_operator", 0, 1, 1), ("grid_block_operator", 64, min(N, 32), 4, 1)],
}


def _check(C, ref, absrow, half):
    tol = 1e-2 if half else 1e-5
    err = np.abs(C.astype(np.float64) - ref)
    bound = tol * np.abs(ref) + (tol if half else 1e-6) * absrow + (2.0 ** -25 if half else 0.0)
    assert (err <= bound).all(), "worst %g at %s" % (err.max(), np.argwhere(err > bound)[0])


@pytest.mark.parametrize("chain", sorted(CHAINS))
@pytest.mark.parametrize("N,half", [(8, False), (32, False), (1, False), (16, True), (128, True), (512, False)])
def test_merge_kernel_on_range_chains(engine, chain, N, half):
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    for name, rows, cols, (r, c) in _mats():
        ip = csr_from_coo(rows, r)
        for ones in (True, False):
            vals = None if ones else synth.synth_values(len(r), seed=9)
            B = np.ones((cols, N), np.float32) if ones else synth.synth_dense(cols, N, seed=9)
            if half and not ones:
                vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
            api.config_set("DENSE_MATRIX_SIZE", N)
            api.config_set("VECTOR_WIDTH", min(N, 32))
            s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
            for op in CHAINS[chain](N):
                s.add_and_run(op[0], *op[1:])
            s.compile()
            d = s.plan_desc()
            assert api.FAMILIES[d.family] == "b200_merge" and "GLOBAL_META_nz_row_indices_0" in s.deps(), (chain, s.deps())
            engine.upload_session(s, half=half, as_id=70)
            plan = engine.plan(s, N, half=half, sub_matrix_id=70)
            assert "k_b200_merge" in plan.kernel_name
            plan.set_dense(B)
            plan.run(1)
            C = plan.output()
            if ones:
                ref = oracle_spmm_f32(rows, N, ip, c, np.ones(len(r), np.float32), B)
                if not half or ref.max() <= 2048:
                    assert np.array_equal(C.astype(np.float32), ref), (name, plan.kernel_name)
            else:
                _check(C, oracle_spmm_f64acc(rows, N, ip, c, vals, B), oracle_spmm_absrow(rows, N, ip, c, vals, B), half)
            plan.run(3)  # C is zeroed per launch: repeated launches do not accumulate
            if ones and (not half or ref.max() <= 2048):
                assert np.array_equal(plan.output().astype(np.float32), ref)


@pytest.mark.parametrize("family,N,half", [("warp_bitmap", 8, False), ("block_bitmap", 16, True), ("warp_bitmap", 64, False), ("thread_total", 8, False), ("warp_total", 256, True),
                                           ("block_total", 4, False)])
def test_merge_kernel_on_driver_chains(engine, family, N, half):
    """the column-direction driver chains (row chunks of <= 64: the row of every chunk gives the row pointer) and the row-direction ones
    (one row per block: the ranges are the rows) switched to the range kernel"""
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    for name, rows, cols, (r, c) in _mats()[:2]:
        vals = synth.synth_values(len(r), seed=4)
        B = synth.synth_dense(cols, N, seed=4)
        if half:
            vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
        s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
        try:
            s.run_family(family, N, 4 if family == "thread_total" else 1, 1)
        except api.GsError as e:
            if e.status == 5:
                continue  # the reference refuses the padding rate on this matrix too
            raise
        engine.upload_session(s, half=half, as_id=71)
        plan = engine.plan(s, N, half=half, kernel="b200_merge", sub_matrix_id=71)
        plan.set_dense(B)
        plan.run(2)
        ip = csr_from_coo(rows, r)
        _check(plan.output(), oracle_spmm_f64acc(rows, N, ip, c, vals, B), oracle_spmm_absrow(rows, N, ip, c, vals, B), half)


def test_total_families_refuse_multi_row_blocks(engine):
    """a row block of 4 rows on the reference-shaped thread_total kernel would store four rows' sum as one row: refused, auto takes the range kernel"""
    from generalsparse_b200 import api, synth
    rows, cols, (r, c) = 700, 450, synth.synth_matrix("ragged", 700, 450, seed=2, mean_len=9.0, empty_frac=0.3)
    api.config_set("DENSE_MATRIX_SIZE", 8)
    api.config_set("VECTOR_WIDTH", 8)
    s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c))
    for op in CHAINS["rowblock4_thread"](8):
        s.add_and_run(op[0], *op[1:])
    s.compile()
    engine.upload_session(s, as_id=72)
    plan = engine.plan(s, 8, sub_matrix_id=72)
    with pytest.raises(api.GsError):
        plan.set_kernel("thread_total")
    with pytest.raises(api.GsError):
        plan.set_kernel("b200_row_smem")
    plan.set_kernel("auto")
    assert "k_b200_merge" in plan.kernel_name


@pytest.mark.parametrize("stage,ticket", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("N,half", [(8, False), (2, False), (32, False), (16, True), (8, True)])
def test_chunk_kernel_modes(engine, monkeypatch, stage, ticket, N, half):
    """the narrow-N variant (one 16-byte chunk of B's row per lane) in its four modes: the CTA's piece of the stream staged through shared memory or
    not; rows shared between segments completed by red.add into a zeroed C, or by tickets (partials summed in stream order by the last segment to
    arrive: no memset, and the result is the same bits on every launch).  Matrices with empty rows in the middle and at the end, sorted chains
    (the row permutation applies to the zero rows too), slices of 5 non-zeros so that every row of the long-row matrix is shared."""
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    monkeypatch.setenv("GS_MERGE_STAGE", str(stage))
    monkeypatch.setenv("GS_MERGE_TICKET", str(ticket))
    rr, rc = synth.synth_matrix("ragged", 700, 450, seed=2, mean_len=9.0, empty_frac=0.3)
    keep = rr < 640  # sixty trailing rows without a non-zero
    lr, lc = synth.synth_matrix("lognormal", 900, 1200, seed=3, mean_len=60.0)
    for name, rows, cols, r, c, cap, ops in (
            ("ragged+tail", 700, 450, rr[keep], rc[keep], 5, CHAINS["merge_path_thread"](N)),
            ("ragged sorted", 700, 450, rr[keep], rc[keep], 32, CHAINS["sort_merge_path_thread"](N)),
            ("long rows", 900, 1200, lr, lc, 16, None),
            ("long rows balanced", 900, 1200, lr, lc, 32, CHAINS["balanced_thread"](N))):
        monkeypatch.setenv("GS_MERGE_CAP", str(cap))
        ip = csr_from_coo(rows, r)
        for ones in (True, False):
            vals = None if ones else synth.synth_values(len(r), seed=9)
            B = np.ones((cols, N), np.float32) if ones else synth.synth_dense(cols, N, seed=9)
            if half and not ones:
                vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
            api.config_set("DENSE_MATRIX_SIZE", N)
            api.config_set("VECTOR_WIDTH", min(N, 32))
            s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
            if ops is None:
                s.run_family("warp_total", N, 1, 1)  # one row per block: the ranges are the rows, cut into slices of <= cap
            else:
                for op in ops:
                    s.add_and_run(op[0], *op[1:])
                s.compile()
            engine.upload_session(s, half=half, as_id=73)
            plan = engine.plan(s, N, half=half, kernel="b200_merge", sub_matrix_id=73)
            kn = plan.kernel_name
            assert "k_b200_merge_chunk" in kn and ("staged," in kn) == bool(stage), kn
            if ticket:
                assert "tickets," in kn, kn  # no row of these matrices is shared by more than 64 slices
            plan.set_dense(B)
            plan.run(1)
            C = plan.output()
            if ones:
                ref = oracle_spmm_f32(rows, N, ip, c, np.ones(len(r), np.float32), B)
                if not half or ref.max() <= 2048:
                    assert np.array_equal(C.astype(np.float32), ref), (name, kn)
            else:
                _check(C, oracle_spmm_f64acc(rows, N, ip, c, vals, B), oracle_spmm_absrow(rows, N, ip, c, vals, B), half)
            plan.run(5)  # back to back (programmatic dependent launch in ticket mode): the counters are back at zero after every launch
            C2 = plan.output()
            if ticket:
                assert np.array_equal(C2.view(np.uint8), C.view(np.uint8)), (name, kn)  # fixed order of the sum: the same bits
            elif ones and (not half or ref.max() <= 2048):
                assert np.array_equal(C2.astype(np.float32), ref)


def test_chunk_kernel_hub_row_keeps_red_add(engine, monkeypatch):
    """a row cut into more than GS_MERGE_TICKET_MAX slices: the plan stays with memset + red.add (the last arriver would sum thousands of partials alone)"""
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    monkeypatch.setenv("GS_MERGE_TICKET", "1")
    rows, cols, N = 300, 5000, 8
    r, c = synth.synth_matrix("uniform", rows, cols, seed=7, density=0.004)
    hub = np.arange(0, cols, 1, dtype=r.dtype)
    keep = r != 17
    r, c = np.concatenate([r[keep], np.full(len(hub), 17, r.dtype)]), np.concatenate([c[keep], hub])
    order = np.lexsort((c, r))
    r, c = r[order], c[order]
    vals, B = synth.synth_values(len(r), seed=1), synth.synth_dense(cols, N, seed=1)
    s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
    s.run_family("warp_total", N, 1, 1)
    engine.upload_session(s, as_id=74)
    plan = engine.plan(s, N, kernel="b200_merge", sub_matrix_id=74)
    assert "k_b200_merge_chunk" in plan.kernel_name and "tickets," not in plan.kernel_name, plan.kernel_name
    plan.set_dense(B)
    plan.run(2)
    ip = csr_from_coo(rows, r)
    _check(plan.output(), oracle_spmm_f64acc(rows, N, ip, c, vals, B), oracle_spmm_absrow(rows, N, ip, c, vals, B), False)


@pytest.mark.parametrize("chain", sorted(CHAINS) + ["warp_bitmap", "thread_total"])
@pytest.mark.parametrize("N,half", [(8, False), (32, True), (128, False), (2, False)])
def test_plan_built_on_device_equals_host_built(engine, monkeypatch, chain, N, half):
    """the range kernel's plan -- segments, their rows, who shares which row, the rows without non-zeros -- built by device passes (the default) against
    the host code (GS_MERGE_DEVICE_PREPARE=0): the same kernel line (segment count, tickets or not), and the same C: bit for bit in ticket mode (the
    order of every sum is fixed by the arrays), within fp32 rounding with red.add"""
    from generalsparse_b200 import api, synth
    for name, rows, cols, (r, c) in _mats():
        vals = synth.synth_values(len(r), seed=9)
        B = synth.synth_dense(cols, N, seed=9)
        if half:
            vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
        api.config_set("DENSE_MATRIX_SIZE", N)
        api.config_set("VECTOR_WIDTH", min(N, 32))
        s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
        try:
            if chain in CHAINS:
                for op in CHAINS[chain](N):
                    s.add_and_run(op[0], *op[1:])
                s.compile()
            else:
                s.run_family(chain, N, 4 if chain == "thread_total" else 1, 1)
        except api.GsError as e:
            if e.status == 5:
                continue  # the reference refuses the padding rate on this matrix
            raise
        engine.upload_session(s, half=half, as_id=75)
        out = {}
        for mode in ("0", "1"):
            monkeypatch.setenv("GS_MERGE_DEVICE_PREPARE", mode)
            plan = engine.plan(s, N, half=half, kernel="b200_merge", sub_matrix_id=75)
            plan.set_dense(B)
            plan.run(2)
            out[mode] = (plan.output(), plan.kernel_name)
            del plan
        (Ch, kh), (Cd, kd) = out["0"], out["1"]
        assert kh == kd, (name, chain, kh, kd)
        if "tickets," in kd:
            assert np.array_equal(Ch.view(np.uint8), Cd.view(np.uint8)), (name, chain, kd)
        else:
            assert np.allclose(Ch.astype(np.float32), Cd.astype(np.float32), rtol=1e-2 if half else 1e-5, atol=1e-2 if half else 1e-5), (name, chain, kd)
