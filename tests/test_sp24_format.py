"""The B200 2:4 image (csrc/host/sp24_format.hpp), on the CPU: build from CSR, expand image + remainder again, compare;
the byte layout of one tile against an independent numpy statement of it.  No GPU involved."""
import ctypes as C

import numpy as np
import pytest


def _roundtrip(rows, cols, indptr, col, vals16, group=1):
    from generalsparse_b200 import _lib
    L = _lib.lib()
    dense = np.zeros((rows, cols), dtype=np.uint16)
    stats = np.zeros(4, dtype=np.uint64)
    rp = np.ascontiguousarray(indptr, dtype=np.uint32)
    cc = np.ascontiguousarray(col, dtype=np.uint32)
    vv = np.ascontiguousarray(vals16.view(np.uint16))
    _lib.check(L.gsh_sp24_roundtrip(rp.ctypes.data_as(C.c_void_p), cc.ctypes.data_as(C.c_void_p), vv.ctypes.data_as(C.c_void_p), rows, cols, group,
                                    dense.ctypes.data_as(C.c_void_p), stats.ctypes.data_as(C.c_void_p)))
    return dense, stats


def _csr_of(mask, seed):
    rng = np.random.default_rng(seed)
    vals = rng.normal(0, 0.02, mask.shape).astype(np.float16)
    vals[vals == 0] = np.float16(0.01)
    dense = np.where(mask, vals, np.float16(0))
    r, c = np.nonzero(mask)
    indptr = np.zeros(mask.shape[0] + 1, dtype=np.int64)
    np.add.at(indptr, r + 1, 1)
    return np.cumsum(indptr), c, dense[r, c], dense


def _mask_24(rows, cols, seed):
    rng = np.random.default_rng(seed)
    w = rng.random((rows, cols // 4, 4))
    keep = np.argsort(-w, axis=2)[:, :, :2]
    mask = np.zeros((rows, cols // 4, 4), dtype=bool)
    np.put_along_axis(mask, keep, True, axis=2)
    return mask.reshape(rows, cols)


@pytest.mark.parametrize("group", [1, 2])
@pytest.mark.parametrize("rows,cols", [(128, 128), (130, 260), (1000, 4096), (5, 8), (300, 128)])
def test_exact_24_has_no_remainder(built_lib, rows, cols, group):
    indptr, col, v, dense = _csr_of(_mask_24(rows, cols, rows + cols), seed=1)
    out, stats = _roundtrip(rows, cols, indptr, col, v, group)
    assert np.array_equal(out, dense.view(np.uint16))
    assert int(stats[1]) == len(col) and int(stats[2]) == 0
    panels = -(-(-(-rows // 128)) // group) * group  # padded to whole units of `group` panels
    assert int(stats[3]) == panels * -(-cols // 128) and int(stats[0]) == int(stats[3]) * 18432


@pytest.mark.parametrize("rows,cols,density,empty", [(130, 70, 0.3, 0), (300, 1000, 0.02, 60), (1000, 4096, 0.3, 0), (257, 129, 0.9, 3), (5, 3, 1.0, 0)])
def test_unstructured_goes_to_image_plus_remainder(built_lib, rows, cols, density, empty):
    rng = np.random.default_rng(rows * cols)
    mask = rng.random((rows, cols)) < density
    if empty:
        mask[rng.choice(rows, empty, replace=False)] = False
    indptr, col, v, dense = _csr_of(mask, seed=2)
    out, stats = _roundtrip(rows, cols, indptr, col, v, 1 + rows % 2)
    assert np.array_equal(out, dense.view(np.uint16))
    # every group of four columns keeps its first two non-zeros, the rest is remainder
    pad = (-cols) % 4
    per_group = np.pad(mask, ((0, 0), (0, pad))).reshape(rows, -1, 4).sum(axis=2)
    assert int(stats[1]) == int(np.minimum(per_group, 2).sum()) and int(stats[2]) == int(np.maximum(per_group - 2, 0).sum())
    if (rows, cols, density) == (1000, 4096, 0.3):  # the C2 density: 7.65 % of the non-zeros expected in the remainder
        assert 0.070 < stats[2] / len(col) < 0.083


def test_chain_padding_is_dropped_and_duplicates_refused(built_lib):
    from generalsparse_b200 import _lib
    indptr, col, v, dense = _csr_of(_mask_24(64, 128, 9), seed=3)
    # pad every row to a multiple of 8 the way the reference does: repeat the last column with value 0
    rows = len(indptr) - 1
    ncol, nval, nptr = [], [], [0]
    for r in range(rows):
        cs, vs = list(col[indptr[r]:indptr[r + 1]]), list(v[indptr[r]:indptr[r + 1]])
        cs += [cs[-1]] * 5
        vs += [np.float16(0)] * 5
        ncol += cs
        nval += vs
        nptr.append(len(ncol))
    out, stats = _roundtrip(rows, 128, np.array(nptr), np.array(ncol), np.array(nval, dtype=np.float16))
    assert np.array_equal(out, dense.view(np.uint16)) and int(stats[2]) == 0
    bad_col, bad_val = np.array([3, 3], dtype=np.uint32), np.array([1.0, 2.0], dtype=np.float16)
    with pytest.raises(_lib.GsError):
        _roundtrip(1, 8, np.array([0, 2]), bad_col, bad_val)


def test_lone_elements_and_metadata_patterns(built_lib):
    """one element per group in each of the four positions, and every pair: the 4-bit codes are the six valid ones"""
    rows, cols = 16, 128
    mask = np.zeros((rows, cols), dtype=bool)
    pats = [(0,), (1,), (2,), (3,), (0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3), ()]
    for r in range(rows):
        for g in range(cols // 4):
            for p in pats[(r * 7 + g) % len(pats)]:
                mask[r, 4 * g + p] = True
    indptr, col, v, dense = _csr_of(mask, seed=4)
    out, stats = _roundtrip(rows, cols, indptr, col, v)
    assert np.array_equal(out, dense.view(np.uint16)) and int(stats[2]) == 0


def _tiled_roundtrip(rows, cols, indptr, col, vals16, row_bytes=32):
    from generalsparse_b200 import _lib
    L = _lib.lib()
    dense = np.zeros((rows, cols), dtype=np.uint16)
    stats = np.zeros(7, dtype=np.uint64)
    rp = np.ascontiguousarray(indptr, dtype=np.uint32)
    cc = np.ascontiguousarray(col, dtype=np.uint32)
    vv = np.ascontiguousarray(vals16.view(np.uint16))
    _lib.check(L.gsh_sp24_tiled_roundtrip(rp.ctypes.data_as(C.c_void_p), cc.ctypes.data_as(C.c_void_p), vv.ctypes.data_as(C.c_void_p), rows, cols, row_bytes,
                                          dense.ctypes.data_as(C.c_void_p), stats.ctypes.data_as(C.c_void_p)))
    return dense, stats


@pytest.mark.parametrize("rows,cols,density,empty", [(130, 70, 0.3, 0), (300, 1000, 0.02, 60), (1000, 4096, 0.3, 0), (257, 129, 0.9, 3), (5, 3, 1.0, 0), (128, 128, 1.0, 0)])
def test_remainder_cut_along_the_tiles(built_lib, rows, cols, density, empty):
    """Every tile image followed by the remainder entries of its window: image + tile-local blocks expand to the matrix, the
    unit is the image + 272 bytes of row starts + the largest block, and the blocks hold what the CSR remainder held."""
    rng = np.random.default_rng(rows * cols + 1)
    mask = rng.random((rows, cols)) < density
    if empty:
        mask[rng.choice(rows, empty, replace=False)] = False
    indptr, col, v, dense = _csr_of(mask, seed=3)
    out, stats = _tiled_roundtrip(rows, cols, indptr, col, v, 32 << (rows % 3))
    assert np.array_equal(out, dense.view(np.uint16))
    _, s0 = _roundtrip(rows, cols, indptr, col, v, 1)
    assert int(stats[1]) == int(s0[1]) and int(stats[2]) == int(s0[2]) and int(stats[3]) == int(s0[3])
    # the largest block, counted independently: third and fourth non-zeros of a 4-column group, per 128 x 128 window
    pr, pc = -(-rows // 128) * 128, -(-cols // 128) * 128
    m = np.zeros((pr, pc), dtype=bool)
    m[:rows, :cols] = mask
    extra = np.maximum(m.reshape(pr, pc // 4, 4).sum(axis=2) - 2, 0)
    per_tile = extra.reshape(pr // 128, 128, pc // 128, 32).sum(axis=(1, 3))
    assert int(stats[5]) == int(per_tile.max())
    assert int(stats[4]) == 18432 + 272 + 4 * (-(-int(per_tile.max()) // 4) * 4) and int(stats[0]) == int(stats[3]) * int(stats[4])


def test_empty_tiles_leave_the_stream(built_lib):
    """a block-sparse matrix: the units kept are the blocks that hold something plus one for every row panel that holds
    nothing, and image + tile-local blocks of the kept units still expand to the matrix"""
    rng = np.random.default_rng(23)
    rows, cols = 1100, 1300  # 9 row panels (the last ragged), 11 k tiles (the last ragged)
    blocks = rng.random((9, 11)) < 0.3
    blocks[2] = False
    blocks[8] = False
    blocks[8, 10] = True
    mask = (rng.random((rows, cols)) < 0.5) & np.kron(blocks, np.ones((128, 128), dtype=bool))[:rows, :cols]
    indptr, col, v, dense = _csr_of(mask, seed=4)
    out, stats = _tiled_roundtrip(rows, cols, indptr, col, v, 64)
    assert np.array_equal(out, dense.view(np.uint16))
    pm = np.zeros((9 * 128, 11 * 128), dtype=bool)
    pm[:rows, :cols] = mask
    held = pm.reshape(9, 128, 11, 128).any(axis=(1, 3))
    assert int(stats[6]) == int(held.sum()) + int((~held.any(axis=1)).sum()) and int(stats[3]) == 99
