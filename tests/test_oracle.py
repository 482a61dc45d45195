"""The oracle itself: oracle/spmm_oracle.c (plain-C restatement of spmm_reference_host,
cuda_code/kernel_lib.hpp:859-921) pinned against (1) the reference program's own golden rule -- all-ones A and B
give C[i][*] == nnz(row i) exactly (code_generator.cc:631-638), (2) the reference's own function compiled from
/root/reference (oracle/_ref/libref_spmm.so) where that build is present, (3) numpy in fp64.  CPU only."""
import ctypes as C
import os

import numpy as np
import pytest

from generalsparse_b200 import synth
from oracle.oracle_check import (csr_from_coo, oracle_lib, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc,
                                             ref_lib, ref_spmm_f32)


@pytest.fixture(scope="module", autouse=True)
def _built():
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(g.ROOT, "oracle", "liboracle_spmm.so")):
        g.build_oracle()


def _case(kind, rows, cols, seed, **p):
    r, c = synth.synth_matrix(kind, rows, cols, seed=seed, **p)
    order = np.lexsort((c, r))
    return r[order], c[order]


CASES = [("uniform", 64, 48, dict(density=0.09)), ("ragged", 200, 150, dict(mean_len=6.0, empty_frac=0.25)),
         ("lognormal", 1500, 1300, dict(mean_len=37.0)), ("powerlaw", 3000, 3000, dict(avg_deg=8.0)), ("uniform", 344, 512, dict(density=0.3))]


@pytest.mark.parametrize("kind,rows,cols,p", CASES)
@pytest.mark.parametrize("N", [1, 8, 32])
def test_golden_rule_all_ones(kind, rows, cols, p, N):
    """the check the reference program runs on every launch: ones x ones == row length, exactly"""
    r, c = _case(kind, rows, cols, 3, **p)
    indptr = csr_from_coo(rows, r)
    out = oracle_spmm_f32(rows, N, indptr, c, np.ones(len(r), np.float32), np.ones((cols, N), np.float32))
    want = np.repeat(np.diff(indptr).astype(np.float32)[:, None], N, axis=1)
    assert np.array_equal(out, want)
    assert oracle_lib().oracle_check_result_f32(C.c_int64(rows), C.c_int64(N), out.ctypes.data_as(C.c_void_p), want.ctypes.data_as(C.c_void_p)) == 0


@pytest.mark.parametrize("kind,rows,cols,p", CASES)
def test_restatement_equals_reference_function(kind, rows, cols, p):
    """bit-exact against the reference's own spmm_reference_host on random values (same loop order, same roundings)"""
    if ref_lib() is None:
        pytest.skip("oracle/_ref/libref_spmm.so not built here (needs /root/reference)")
    r, c = _case(kind, rows, cols, 5, **p)
    indptr = csr_from_coo(rows, r)
    vals = synth.synth_values(len(r), seed=5)
    for N in (4, 8):
        B = synth.synth_dense(cols, N, seed=6)
        mine = oracle_spmm_f32(rows, N, indptr, c, vals, B)
        ref = ref_spmm_f32(rows, N, cols, indptr, c, vals, B)
        assert np.array_equal(mine, ref)
        assert np.array_equal(oracle_spmm_f32(rows, N, indptr, c, vals, B, mt=True), ref)
        assert np.array_equal(ref_spmm_f32(rows, N, cols, indptr, c, vals, B, threads=4), ref)
        # the reference's own check_result (1 = passes its exact compare)
        assert ref_lib().ref_check_result_f32(rows, N, mine.ctypes.data_as(C.c_void_p), ref.ctypes.data_as(C.c_void_p)) == 1


def test_f64_anchor_and_absrow_match_numpy():
    rows, cols, N = 300, 257, 8
    r, c = _case("ragged", rows, cols, 9, mean_len=12.0, empty_frac=0.1)
    indptr = csr_from_coo(rows, r)
    vals = synth.synth_values(len(r), seed=1)
    B = synth.synth_dense(cols, N, seed=2)
    want = np.zeros((rows, N))
    absw = np.zeros((rows, N))
    np.add.at(want, r, vals.astype(np.float64)[:, None] * B.astype(np.float64)[c])
    np.add.at(absw, r, np.abs(vals.astype(np.float64)[:, None] * B.astype(np.float64)[c]))
    got = oracle_spmm_f64acc(rows, N, indptr, c, vals, B)
    assert np.allclose(got, want, rtol=1e-13, atol=1e-15)
    assert np.allclose(oracle_spmm_absrow(rows, N, indptr, c, vals, B), absw, rtol=1e-13, atol=1e-15)
    # the fp32 restatement stays within the tolerance north_star states for fp32 (1e-5 of the row's L1 mass)
    f32 = oracle_spmm_f32(rows, N, indptr, c, vals, B).astype(np.float64)
    assert (np.abs(f32 - want) <= 1e-5 * absw + 1e-30).all()


def test_empty_matrix_and_empty_rows():
    N = 4
    out = oracle_spmm_f32(5, N, np.zeros(6, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32), np.ones((3, N), np.float32))
    assert out.shape == (5, N) and not out.any()


def test_check_result_flags_a_difference():
    a = np.ones((4, 4), np.float32)
    b = a.copy()
    b[2, 1] = 1.0000001
    n = oracle_lib().oracle_check_result_f32(C.c_int64(4), C.c_int64(4), a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p))
    assert n == 1
