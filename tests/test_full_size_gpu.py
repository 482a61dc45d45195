"""Parity at the REAL sizes of BASELINE.json's configs (round 1 tested toy sizes only), through the C ABI, on whatever kernel
`auto` picks (gs_plan_set_kernel(GS_KERNEL_AUTO)) and on named kernels where the config names them:

  C2  11 008 x 4 096, 70 % unstructured, HALF, N = 8 / 16 / 32
  C3  the same weight 90 % unstructured and 2:4-pruned, HALF, N = 64 / 128 / 256
  C4  a 1/10-scale block of the power-law matrix (1 M rows, ~20 M nnz, N = 32, fp32) that CONTAINS its hub row
  C5  65 536 x 65 536 uniform 1 % and banded (half bandwidth 64), the widths each kernel accepts

Gates: reference mode (all-ones A and B) exact where C's type holds the row sums exactly; random values
max |C - C64| <= tol * |C64| + 1e-6 * sum|a*b| (+ 2^-25 for fp16 C) with tol 1e-5 (fp32) / 1e-2 (HALF, fp32 accumulate)
against the fp64-accumulate oracle on the same (fp16-rounded) inputs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(built_lib):
    from generalsparse_b200 import api
    return api.Engine(0)


def _indptr(rows, r):
    ip = np.zeros(rows + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=rows), out=ip[1:])
    return ip


def _gate(C, ref64, absrow, half):
    C = C.astype(np.float64)
    err = np.abs(C - ref64)
    tol = 1e-2 if half else 1e-5
    bound = tol * np.abs(ref64) + 1e-6 * absrow + (2.0 ** -25 if half else 0.0)
    bad = err > bound
    assert not bad.any(), "%d of %d outside the gate; worst err %g (ref %g, row L1 %g) at %s" % (
        bad.sum(), bad.size, err[bad].max(), ref64[bad][np.argmax(err[bad])], absrow[bad][np.argmax(err[bad])], np.argwhere(bad)[0])
    return float((err / np.maximum(np.abs(ref64), 1e-30))[np.abs(ref64) > 1e-3 * absrow].max()) if (np.abs(ref64) > 1e-3 * absrow).any() else 0.0


def _spmm(engine, chain, rows, cols, r, c, vals, B, N, S, half, kernel, as_id):
    from generalsparse_b200 import api
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family(chain, N, S, 1)
    engine.upload_session(s, half=half, as_id=as_id)
    plan = engine.plan(s, N, half=half, kernel=kernel, sub_matrix_id=as_id)
    plan.set_dense(B)
    plan.run(1)
    return plan.output(), plan


def _random_case(engine, chain, rows, cols, r, c, N, S, half, kernel, seed, as_id=60):
    from generalsparse_b200 import synth
    from oracle.oracle_check import oracle_spmm_absrow, oracle_spmm_f64acc
    vals = synth.synth_values(len(r), seed=seed)
    B = synth.synth_dense(cols, N, seed=seed)
    if half:
        vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
    C, plan = _spmm(engine, chain, rows, cols, r, c, vals, B, N, S, half, kernel, as_id)
    ip = _indptr(rows, r)
    ci = c.astype(np.int32)
    worst = _gate(C, oracle_spmm_f64acc(rows, N, ip, ci, vals, B), oracle_spmm_absrow(rows, N, ip, ci, vals, B), half)
    assert plan.algorithmic_bytes > 0 and plan.format_bytes > 0
    return plan, worst


@pytest.fixture(scope="module")
def ffn_patterns():
    from generalsparse_b200 import synth
    rows, cols = 11008, 4096
    return {"70": synth.synth_matrix("uniform", rows, cols, seed=2, density=0.30), "90": synth.synth_matrix("uniform", rows, cols, seed=3, density=0.10),
            "nm24": synth.synth_matrix("nm24", rows, cols, seed=3)}


@pytest.mark.parametrize("N", [8, 16, 32])
def test_c2_full_size_auto(engine, ffn_patterns, N):
    """C2: the 70 %-sparse FFN weight at its real size on the kernel the library picks; SURVEY 8(d) bytes for N = 8"""
    r, c = ffn_patterns["70"]
    plan, worst = _random_case(engine, "thread_total", 11008, 4096, r, c, N, 8, True, "auto", seed=20 + N)
    assert "k_b200_" in plan.kernel_name, plan.kernel_name
    if N == 8:  # nnz_stored * 4 + 2 * 11 009 * 4 + 11 008 * 4 + K * N * 2 + M * N * 2, whatever the kernel streams
        stored = plan.desc.stored_nnz
        assert plan.algorithmic_bytes == stored * 4 + 2 * 11009 * 4 + 11008 * 4 + 4096 * 8 * 2 + 11008 * 8 * 2
    # all-ones mode at full size: exact (row sums ~1 229 <= 2 048 are exact fp16 integers)
    B1 = np.ones((4096, N), dtype=np.float32)
    C, plan = _spmm(engine, "thread_total", 11008, 4096, r, c, None, B1, N, 8, True, "auto", 61)
    assert np.array_equal(C[:, 0].astype(np.int64), np.bincount(r, minlength=11008)) and (C == C[:, :1]).all()


@pytest.mark.parametrize("N", [64, 128, 256])
@pytest.mark.parametrize("pattern", ["90", "nm24"])
def test_c3_full_size_auto(engine, ffn_patterns, pattern, N):
    """C3: the 90 %-sparse and the 2:4-pruned weight at N = 64 ... 256 on the kernel the library picks"""
    r, c = ffn_patterns[pattern]
    plan, worst = _random_case(engine, "thread_total", 11008, 4096, r, c, N, 8, True, "auto", seed=30 + N)
    assert "k_b200_" in plan.kernel_name, plan.kernel_name
    if pattern == "nm24":
        assert "sp24" in plan.kernel_name and "remainder" not in plan.kernel_name, plan.kernel_name


@pytest.mark.parametrize("kernel", ["b200_tcgen05", "b200_sp24", "b200_row_smem"])
def test_c2_full_size_named_kernels(engine, ffn_patterns, kernel):
    r, c = ffn_patterns["70"]
    _random_case(engine, "thread_total", 11008, 4096, r, c, 32 if kernel != "b200_row_smem" else 8, 8, True, kernel, seed=44)


def test_c4_tenth_scale_block_with_hub_row(engine):
    """C4 at 1/10 scale: rows [lo, hi) of the 10 M-row power-law matrix chosen so that the block CONTAINS the hub row (the
    longest row: 2.47 M non-zeros capped to the block's width), N = 32, fp32, nnz-direction bit-map chain on the B200 nnz
    kernel and on the reference-shaped thread_bit_map family; ones exact, values within 1e-5"""
    from generalsparse_b200 import synth
    from oracle.oracle_check import oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    rows10, cols = 10_000_000, 1_000_000
    lengths = synth.powerlaw_lengths(rows10, seed=4, avg_deg=20.0)
    hub = int(np.argmax(lengths))
    lo = max(0, min(hub - 500_000, rows10 - 1_000_000))
    hi = lo + 1_000_000
    assert lo <= hub < hi
    r, c = synth.powerlaw_block(rows10, cols, lengths, lo, hi, seed=4)
    rows = hi - lo
    per_row = np.bincount(r, minlength=rows)
    assert per_row.max() > 100_000 and per_row.argmax() == hub - lo, (per_row.max(), hub - lo)
    N = 32
    ip, ci = _indptr(rows, r), c.astype(np.int32)
    for kernel in ("b200_nnz", None):
        B1 = np.ones((cols, N), dtype=np.float32)
        C, plan = _spmm(engine, "warp_segment" if kernel else "thread_bit_map", rows, cols, r, c, None, B1, N, 1, False, kernel, 62)
        assert np.array_equal(C[:, 0].astype(np.int64), per_row) and (C == C[:, :1]).all(), plan.kernel_name
        vals = synth.synth_values(len(r), seed=7, scale=1.0)
        B = synth.synth_dense(cols, N, seed=7)
        C, plan = _spmm(engine, "warp_segment" if kernel else "thread_bit_map", rows, cols, r, c, vals, B, N, 1, False, kernel, 62)
        _gate(C, oracle_spmm_f64acc(rows, N, ip, ci, vals, B), oracle_spmm_absrow(rows, N, ip, ci, vals, B), False)
    # the hub row itself against the reference's own fp32 loop order (one row: cheap)
    assert abs(C[hub - lo]).max() > 0


@pytest.fixture(scope="module")
def c5_patterns():
    from generalsparse_b200 import synth
    n = 65536
    return {"uniform": synth.synth_matrix("uniform", n, n, seed=5, density=0.01), "banded": synth.synth_matrix("banded", n, n, seed=5, half_bandwidth=64)}


@pytest.mark.parametrize("pattern,chain,kernel,N,half", [
    ("uniform", "thread_total", "auto", 8, True), ("uniform", "thread_total", "auto", 32, True), ("uniform", "thread_total", "auto", 64, True),
    ("uniform", "thread_total", "auto", 8, False), ("uniform", "warp_total", None, 1, False), ("uniform", "thread_total", None, 512, False),
    ("uniform", "warp_segment", "auto", 32, False),
    ("banded", "thread_total", "auto", 8, True), ("banded", "thread_total", "auto", 64, True), ("banded", "thread_total", "auto", 256, True),
    ("banded", "warp_total", None, 2, False), ("banded", "thread_total", None, 512, True), ("banded", "thread_bit_map", "auto", 64, False),
], ids=lambda v: str(v))
def test_c5_full_size(engine, c5_patterns, pattern, chain, kernel, N, half):
    """C5 at the size SURVEY.md 8(d) names (65 536^2; round 1 swept 16 384^2): uniform 1 % (43 M nnz) and banded (8.4 M nnz)"""
    r, c = c5_patterns[pattern]
    S = 8 if chain == "thread_total" and kernel == "auto" else 4 if chain == "thread_total" else 1
    _random_case(engine, chain, 65536, 65536, r, c, N, S, half, kernel, seed=50 + N, as_id=63)
