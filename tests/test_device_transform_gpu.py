"""the transform steps of all seven driver chains run as CUDA passes (csrc/device/transform.cu: binary searches, block
scans, a stable LSD radix sort, warp ballots) instead of the host pipeline.  The bar is the host pipeline itself, which tests/test_golden.py pins
byte for byte to the reference's dumps: every device array the chain registers must come out with the same dtype and the same bytes, the plan
descriptor must be the one compile() gives, and a plan on the device-built arrays must pick the same kernel and produce the same C."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(built_lib):
    from generalsparse_b200 import api
    return api.Engine(0)


def _mats():
    from generalsparse_b200 import synth
    rr, rc = synth.synth_matrix("ragged", 700, 450, seed=2, mean_len=9.0, empty_frac=0.3)
    keep = rr < 640  # sixty trailing rows without a non-zero
    wr, wc = synth.synth_matrix("uniform", 300, 70000, seed=4, density=0.0008)  # column ids beyond 16 bits
    return [("ragged+tail", 700, 450, rr[keep], rc[keep]),
            ("powerlaw", 3000, 3000) + tuple(synth.synth_matrix("powerlaw", 3000, 3000, seed=5, avg_deg=20.0, alpha=2.1)),
            ("lognormal", 5000, 4096) + tuple(synth.synth_matrix("lognormal", 5000, 4096, seed=3, mean_len=37.0)),
            ("wide", 300, 70000, wr, wc),
            ("tiny", 64, 48) + tuple(synth.synth_matrix("uniform", 64, 48, seed=1, density=0.085))]


def _desc_fields(d, strip_id):
    deps = [bytes(d.deps[i]).split(b"\0")[0].decode() for i in range(d.n_deps)]
    deps = [k[: k.rfind("_")] for k in deps] if strip_id else deps
    return dict(family=d.family, S=d.sparse_coarsen, C=d.coarsen, vw=d.vector_width, warp=d.need_warp_reduction, tfr=d.thread_for_row, perm=d.has_row_permutation,
                grid=tuple(d.grid), block=tuple(d.block), rows=d.rows, cols=d.cols, nnz=d.origin_nnz, stored=d.stored_nnz, deps=deps)


@pytest.mark.parametrize("family,S", [("thread_total", 2), ("thread_total", 8), ("warp_total", 1), ("thread_bit_map", 1), ("warp_segment", 1), ("block_total", 1), ("warp_total", 4), ("warp_bitmap", 1), ("block_bitmap", 4)])
@pytest.mark.parametrize("N,half", [(8, False), (32, True), (1, False)])
def test_device_chain_equals_host_chain(engine, family, S, N, half):
    from generalsparse_b200 import api, synth
    for name, rows, cols, r, c in _mats():
        vals = synth.synth_values(len(r), seed=9)
        s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
        try:
            s.run_family(family, N, S, 1)
            host_err = None
        except api.GsError as e:
            host_err = e.status
        if host_err is not None:  # the reference refuses (padding rate): the device chain refuses with the same status
            with pytest.raises(api.GsError) as ei:
                engine.transform_coo(family, rows, cols, r, c, vals, N, S, 1, half=half, as_id=81)
            assert ei.value.status == host_err, (name, family, ei.value)
            continue
        engine.upload_session(s, half=half, as_id=80)
        hd = s.plan_desc()
        dd, ms = engine.transform_coo(family, rows, cols, r, c, vals, N, S, 1, half=half, as_id=81)
        assert ms > 0
        assert _desc_fields(hd, True) == _desc_fields(dd, True), (name, family)
        assert dd.sub_matrix_id == 81
        keys = [k[: k.rfind("_")] for k in s.deps()] + ["GLOBAL_META_nz_row_indices"]
        for k in keys:
            a, b = engine.download_array(k + "_80"), engine.download_array(k + "_81")
            assert a.dtype == b.dtype and a.shape == b.shape, (name, family, k, a.dtype, b.dtype, a.shape, b.shape)
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), (name, family, k, np.flatnonzero(a != b)[:5])
        # the same kernel, the same C (values as uploaded; the reference-shaped family and the library's own choice)
        B = synth.synth_dense(cols, N, seed=9)
        if half:
            B = B.astype(np.float16).astype(np.float32)
        for kernel in (None, "auto"):
            ph = engine.plan(s, N, half=half, kernel=kernel, sub_matrix_id=80)
            pd = engine.plan(dd, N, half=half, kernel=kernel)
            assert ph.kernel_name == pd.kernel_name, (name, family, kernel)
            ph.set_dense(B), pd.set_dense(B)
            ph.run(1), pd.run(1)
            Ch, Cd = ph.output().astype(np.float32), pd.output().astype(np.float32)
            if family in ("thread_bit_map", "warp_segment", "warp_bitmap", "block_bitmap") or "merge" in pd.kernel_name or "nnz" in pd.kernel_name:  # atomics: the order of the adds is free
                assert np.allclose(Ch, Cd, rtol=1e-2 if half else 1e-5, atol=1e-2 if half else 1e-5), (name, family, kernel)
            else:
                assert np.array_equal(Ch, Cd), (name, family, kernel)
            del ph, pd


def test_device_chain_refusals(engine):
    from generalsparse_b200 import api
    r = np.array([0, 1, 2, 3], np.uint32)
    c = np.array([1, 0, 3, 2], np.uint32)
    with pytest.raises(api.GsError) as ei:  # every row grows from 1 to 8 entries: 8x >= PADDING_RATE_UP_BOUND
        engine.transform_coo("thread_total", 4, 4, r, c, None, 8, 8, 1, as_id=82)
    assert ei.value.status == 5
    with pytest.raises(api.GsError):  # multiple_of_each_row_size must be >= 2 (the reference asserts)
        engine.transform_coo("thread_total", 4, 4, r, c, None, 8, 1, 1, as_id=82)
    with pytest.raises(api.GsError):  # not row-major
        engine.transform_coo("warp_total", 4, 4, r[::-1].copy(), c, None, 8, 1, 1, as_id=82)
    with pytest.raises(api.GsError):  # a column id outside the matrix
        engine.transform_coo("warp_total", 4, 3, r, c, None, 8, 1, 1, as_id=82)
    with pytest.raises(api.GsError) as ei:
        engine.transform_coo("b200_merge", 4, 4, r, c, None, 8, 1, 1, as_id=82)
    assert ei.value.status == 3


def test_device_chain_sort_is_stable_at_scale(engine):
    """300 000 rows whose lengths need two radix passes (up to 700): the permutation is the reference's counting sort, longest first, ties in row order"""
    from generalsparse_b200 import api
    rng = np.random.default_rng(11)
    rows, cols = 300_000, 5000
    lens = rng.integers(0, 12, rows)
    lens[rng.integers(0, rows, 200)] = rng.integers(256, 700, 200)
    r = np.repeat(np.arange(rows, dtype=np.uint32), lens)
    c = rng.integers(0, cols, len(r)).astype(np.uint32)
    dd, ms = engine.transform_coo("thread_total", rows, cols, r, c, None, 8, 2, 1, as_id=83)
    order = engine.download_array("GLOBAL_META_original_nz_row_indices_83")
    expect = np.argsort(-lens, kind="stable").astype(np.uint32)
    assert np.array_equal(order, expect)
    kept = int((lens > 0).sum())
    fn = engine.download_array("THREAD_META_first_nz_indices_83")
    plen = (lens[expect[:kept]] + 1) // 2 * 2
    assert np.array_equal(fn, np.concatenate([[0], np.cumsum(plen)]).astype(np.uint32))
    assert dd.stored_nnz == plen.sum()


@pytest.mark.parametrize("level,work,sort", [("thread", 33, False), ("thread", 7, True), ("warp", 300, False), ("tblock", 4096, False), ("thread", 1, False)])
@pytest.mark.parametrize("N,half", [(8, False), (32, True)])
def test_device_merge_path_chain_equals_host_chain(engine, level, work, sort, N, half):
    """[sort_operator +] merge_path_{thread,warp,tblock}_operator(work) + the level's total reduction: the partition by one binary search per part on the
    device against the host's two-pointer walk (itself pinned to the reference's rows x parts loop by the goldens): same arrays, same descriptor, same kernel, same C"""
    from generalsparse_b200 import api, synth
    red = {"thread": ("thread_total_reduce_operator", 0, 1, 1), "warp": ("warp_total_reduce_operator", 1), "tblock": ("tblock_total_reduce_operator", 1)}[level]
    bx, by = min(N, 32), 4
    for name, rows, cols, r, c in _mats():
        vals = synth.synth_values(len(r), seed=9)
        api.config_set("DENSE_MATRIX_SIZE", N)
        api.config_set("VECTOR_WIDTH", min(N, 32))
        s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals))
        for op in ([("sort_operator",)] if sort else []) + [("merge_path_%s_operator" % level, work), red, ("grid_block_operator", 64, bx, by, 1)]:
            s.add_and_run(op[0], *op[1:])
        s.compile()
        hd = s.plan_desc()
        engine.upload_session(s, half=half, as_id=84)
        dd = engine.transform_coo_merge_path(level, work, rows, cols, r, c, vals, N, 1, min(N, 32), tuple(hd.grid), tuple(hd.block), sort=sort, half=half, as_id=85)
        assert _desc_fields(hd, True) == _desc_fields(dd, True), (name, level)
        for k in [k[: k.rfind("_")] for k in s.deps()]:
            a, b = engine.download_array(k + "_84"), engine.download_array(k + "_85")
            assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), (name, level, k)
        B = synth.synth_dense(cols, N, seed=9)
        if half:
            B = B.astype(np.float16).astype(np.float32)
        ph, pd = engine.plan(s, N, half=half, sub_matrix_id=84), engine.plan(dd, N, half=half)
        assert ph.kernel_name == pd.kernel_name and "k_b200_merge" in pd.kernel_name, (name, level)
        ph.set_dense(B), pd.set_dense(B)
        ph.run(1), pd.run(1)
        if "tickets," in pd.kernel_name:
            assert np.array_equal(ph.output().view(np.uint8), pd.output().view(np.uint8)), (name, level)
        else:
            tol = 1e-2 if half else 1e-5  # red.add: the order of the adds is free (fp16 partials when C is HALF)
            assert np.allclose(ph.output().astype(np.float32), pd.output().astype(np.float32), rtol=tol, atol=tol), (name, level)
        del ph, pd
