"""The B200 tile stream (csrc/host/tile_format.hpp), on the CPU: build from CSR, expand again, compare; and the effect of
the bank ordering on the shared-memory wavefronts the decompression stores need.  No GPU involved."""
import ctypes as C

import numpy as np
import pytest


def _roundtrip(rows, cols, indptr, col, vals16, pass_warps):
    from generalsparse_b200 import _lib
    L = _lib.lib()
    dense = np.zeros((rows, cols), dtype=np.uint16)
    stats = np.zeros(5, dtype=np.uint64)
    rp = np.ascontiguousarray(indptr, dtype=np.uint32)
    cc = np.ascontiguousarray(col, dtype=np.uint32)
    vv = np.ascontiguousarray(vals16.view(np.uint16))
    _lib.check(L.gsh_tile_stream_roundtrip(rp.ctypes.data_as(C.c_void_p), cc.ctypes.data_as(C.c_void_p), vv.ctypes.data_as(C.c_void_p), rows, cols, pass_warps,
                                           dense.ctypes.data_as(C.c_void_p), stats.ctypes.data_as(C.c_void_p)))
    return dense, stats


def _csr(rows, cols, density, seed, empty_rows=0):
    rng = np.random.default_rng(seed)
    mask = rng.random((rows, cols)) < density
    if empty_rows:
        mask[rng.choice(rows, empty_rows, replace=False)] = False
    vals = rng.normal(0, 0.02, (rows, cols)).astype(np.float16)
    vals[vals == 0] = np.float16(0.01)
    dense = np.where(mask, vals, np.float16(0))
    r, c = np.nonzero(mask)
    indptr = np.zeros(rows + 1, dtype=np.int64)
    np.add.at(indptr, r + 1, 1)
    return np.cumsum(indptr), c, dense[r, c], dense


@pytest.mark.parametrize("rows,cols,density,empty", [(130, 70, 0.3, 0), (300, 1000, 0.02, 60), (1000, 4096, 0.1, 0), (257, 129, 0.9, 3), (5, 3, 1.0, 0)])
@pytest.mark.parametrize("pass_warps", [0, 10])
def test_roundtrip(built_lib, rows, cols, density, empty, pass_warps):
    indptr, col, v, dense = _csr(rows, cols, density, seed=rows + cols, empty_rows=empty)
    out, stats = _roundtrip(rows, cols, indptr, col, v, pass_warps)
    assert np.array_equal(out, dense.view(np.uint16))
    pairs, nnz, tiles = int(stats[0]), int(stats[1]), int(stats[2])
    assert nnz == len(col) and pairs % 8 == 0 and pairs - nnz < 8 * tiles
    assert tiles == -(-rows // 128) * -(-cols // 64)


def test_chain_padding_is_dropped(built_lib):
    """zeros a chain's column padding added (modify_*_by_col_pad_in_sub_matrix.cc) carry nothing: they are not stored"""
    indptr, col, v, dense = _csr(200, 256, 0.2, seed=3)
    # pad every row to a multiple of 8 the way the reference does: repeat the last column with value 0
    ip2, c2, v2 = [0], [], []
    for i in range(200):
        cs, vs = list(col[indptr[i]:indptr[i + 1]]), list(v[indptr[i]:indptr[i + 1]])
        while cs and len(cs) % 8:
            cs.append(cs[-1]), vs.append(np.float16(0))
        c2 += cs
        v2 += vs
        ip2.append(len(c2))
    out, stats = _roundtrip(200, 256, np.array(ip2), np.array(c2), np.array(v2, dtype=np.float16), 10)
    assert np.array_equal(out, dense.view(np.uint16)) and int(stats[1]) == len(col)


def test_bank_order_cuts_store_conflicts(built_lib):
    """on the bench workload's density the ordered runs need < 1.25 wavefronts per warp-wide store; row-major runs need > 2"""
    indptr, col, v, _ = _csr(1024, 2048, 0.3, seed=9)
    _, plain = _roundtrip(1024, 2048, indptr, col, v, 0)
    _, ordered = _roundtrip(1024, 2048, indptr, col, v, 10)
    assert plain[3] == ordered[3]
    assert ordered[4] / ordered[3] < 1.25 < 2.0 < plain[4] / plain[3]


def test_bad_column_is_a_status(built_lib):
    from generalsparse_b200 import _lib
    with pytest.raises(_lib.GsError):
        _roundtrip(4, 8, np.array([0, 1, 1, 1, 1]), np.array([9]), np.array([1.0], dtype=np.float16), 10)
