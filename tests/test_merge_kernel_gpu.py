"""The B200 range kernel (GS_FAMILY_B200_MERGE): the device consumer of merge-path, nnz-balanced, column-direction and multi-row-block
chains (round 1 produced their host arrays bit-exactly but nothing on the device read them).  Parity with the oracle through the C ABI:
all-ones exact, random values within SURVEY.md 8(d)'s gate; fp16 C is completed with red.add at segment edges (each partial rounded to
fp16), hence the row-L1 floor at the tolerance for HALF, as for the other atomic kernels."""
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
