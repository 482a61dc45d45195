"""Deterministic synthetic sparse matrices of the shapes BASELINE.json names.

The reference's example matrices are missing blobs, so every
config runs on a seeded stand-in.  All generators return a row-major-sorted COO pattern
(row ascending, col ascending inside a row, no duplicates) as int64 numpy arrays -- the
order the reference loader demands (struct.cc:125).
"""
import numpy as np


def _sorted_unique(key):
    """np.unique for int64 keys by sort + neighbour compare (numpy's own is 50x slower on millions of keys)"""
    key = np.sort(key)
    if len(key) == 0:
        return key
    keep = np.empty(len(key), dtype=bool)
    keep[0] = True
    np.not_equal(key[1:], key[:-1], out=keep[1:])
    return key[keep]


def _rows_from_lengths(lengths):
    return np.repeat(np.arange(len(lengths), dtype=np.int64), lengths)


def _cols_for_lengths(rng, lengths, cols, weights=None):
    """Distinct sorted column ids per row, vectorised: sample keys then dedupe per row."""
    lengths = np.minimum(lengths, cols).astype(np.int64)
    total = int(lengths.sum())
    r = _rows_from_lengths(lengths)
    if weights is None:
        c = rng.integers(0, cols, size=total, dtype=np.int64)
    else:
        cdf = np.cumsum(weights)
        cdf /= cdf[-1]
        c = np.searchsorted(cdf, rng.random(total)).astype(np.int64)
        c = np.minimum(c, cols - 1)
    key = r * cols + c
    key = _sorted_unique(key)  # sorted, duplicates dropped (row lengths shrink slightly)
    return key // cols, key % cols


def synth_matrix(kind, rows, cols, seed=0, **p):
    """Return (row, col) int64 arrays of a seeded synthetic pattern.

    kinds: uniform(density) | ragged(mean_len, empty_frac) | lognormal(mean_len) |
           powerlaw(avg_deg, alpha) | banded(half_bandwidth) | nm24()
    """
    rng = np.random.default_rng(0x6753 + seed)
    if kind == "uniform":
        density = p["density"]
        if rows * cols <= (1 << 27):
            mask = rng.random((rows, cols)) < density
            r, c = np.nonzero(mask)
            return r.astype(np.int64), c.astype(np.int64)
        lengths = rng.binomial(cols, density, size=rows)
        return _cols_for_lengths(rng, lengths, cols)
    if kind == "ragged":
        lengths = rng.geometric(1.0 / p["mean_len"], size=rows)
        lengths[rng.random(rows) < p.get("empty_frac", 0.1)] = 0
        lengths[rows - max(1, rows // 50):] = 0  # trailing empty rows (remove_empty_row_in_end path)
        if rows > 8:
            lengths[3] = min(cols, 5 * int(p["mean_len"]) + 70)  # one long row (> 64: several col-dir chunks)
        return _cols_for_lengths(rng, lengths, cols)
    if kind == "lognormal":
        sigma = 0.9
        mu = np.log(p["mean_len"]) - sigma * sigma / 2
        lengths = np.maximum(1, rng.lognormal(mu, sigma, size=rows)).astype(np.int64)
        return _cols_for_lengths(rng, lengths, cols)
    if kind == "powerlaw":
        alpha = p.get("alpha", 2.1)
        u = rng.random(rows)
        deg = np.floor((1.0 - u) ** (-1.0 / (alpha - 1.0))).astype(np.float64)
        deg = np.minimum(deg, p.get("cap", max(8, cols // 10)))
        deg *= p["avg_deg"] * rows / deg.sum()
        lengths = np.maximum(0, np.round(deg)).astype(np.int64)
        w = (np.arange(1, cols + 1, dtype=np.float64)) ** (-0.6)
        w = w[rng.permutation(cols)]
        return _cols_for_lengths(rng, lengths, cols, weights=w)
    if kind == "banded":
        hb = p["half_bandwidth"]
        offs = np.arange(-hb, hb + 1, dtype=np.int64)
        r = np.repeat(np.arange(rows, dtype=np.int64), len(offs))
        c = r + np.tile(offs, rows)
        ok = (c >= 0) & (c < cols)
        return r[ok], c[ok]
    if kind == "nm24":
        # exactly 2 of every 4 consecutive columns kept (magnitude-chosen from random weights)
        assert cols % 4 == 0
        w = rng.random((rows, cols // 4, 4))
        keep = np.argsort(-w, axis=2)[:, :, :2]
        keep.sort(axis=2)
        base = (np.arange(cols // 4, dtype=np.int64) * 4)[None, :, None]
        c = (keep + base).reshape(rows, -1)
        r = np.repeat(np.arange(rows, dtype=np.int64), c.shape[1])
        return r, c.reshape(-1).astype(np.int64)
    raise ValueError("unknown synthetic kind %r" % kind)


def powerlaw_lengths(rows, seed=0, avg_deg=20.0, alpha=2.1, cap=1000000):
    """Row degrees of the C4 stand-in (10 M-row power-law matrix): Zipf-like tail with exponent alpha, capped, scaled to
    rows * avg_deg non-zeros in total.  Cheap and global: every rank computes the same array and the same row cuts."""
    rng = np.random.default_rng(0x6753 + seed)
    u = rng.random(rows)
    deg = np.floor((1.0 - u) ** (-1.0 / (alpha - 1.0)))
    deg = np.minimum(deg, cap)
    deg *= avg_deg * rows / deg.sum()
    return np.maximum(0, np.round(deg)).astype(np.int64)


def powerlaw_block(rows, cols, lengths, lo, hi, seed=0, chunk=1 << 16):
    """Rows [lo, hi) of the C4 stand-in as (row - lo, col), row-major sorted.  Column ids follow a power-law popularity
    (weight rank^-0.6 over a fixed permutation of the columns).  Rows are generated in fixed chunks with one generator
    per chunk, so the matrix does not depend on how it is cut over the ranks."""
    # popularity rank^-0.6: the CDF is ~ rank^0.4, so rank = cols * u^2.5 by inversion; a fixed permutation scatters the
    # popular columns over the id range
    perm = np.random.default_rng(0x6753 + seed).permutation(cols).astype(np.int64)
    out_r, out_c = [], []
    for k in range(lo // chunk, (hi + chunk - 1) // chunk):
        a, b = k * chunk, min(rows, (k + 1) * chunk)
        rng = np.random.default_rng([0x6753 + seed, k])
        ln = np.minimum(lengths[a:b], cols)
        r = np.repeat(np.arange(a, b, dtype=np.int64), ln)
        c = perm[np.minimum((rng.random(int(ln.sum()), dtype=np.float32).astype(np.float64) ** 2.5 * cols).astype(np.int64), cols - 1)]
        key = _sorted_unique(r * cols + c)  # sorted; duplicate draws inside a row are dropped
        r, c = key // cols, key % cols
        sel = (r >= lo) & (r < hi)
        out_r.append(r[sel] - lo), out_c.append(c[sel])
    if not out_r:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    return np.concatenate(out_r), np.concatenate(out_c)


def synth_values(nnz, seed=0, scale=0.02, dtype=np.float32):
    """Pruned-weight-like values N(0, scale); reference mode uses all ones instead."""
    rng = np.random.default_rng(0x7A11 + seed)
    return (rng.standard_normal(nnz) * scale).astype(dtype)


def synth_dense(k, n, seed=0, dtype=np.float32):
    rng = np.random.default_rng(0xB0B + seed)
    return rng.uniform(-1.0, 1.0, size=(k, n)).astype(dtype)


def exact_lengths_matrix(rng, lengths, cols):
    """Distinct uniform column ids, EXACTLY lengths[i] of them in row i (row-major sorted): candidates are over-drawn,
    de-duplicated keeping the first draw, and every row keeps its first lengths[i] distinct ones."""
    lengths = np.minimum(lengths, cols).astype(np.int64)
    draw = lengths + lengths // 8 + 4
    r = _rows_from_lengths(draw)
    c = rng.integers(0, cols, size=int(draw.sum()), dtype=np.int64)
    key = r * cols + c
    _, first = np.unique(key, return_index=True)
    first.sort()  # distinct (row, col) pairs in draw order; rows stay grouped because r is non-decreasing
    r, c = r[first], c[first]
    start = np.zeros(len(lengths) + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=len(lengths)), out=start[1:])
    rank = np.arange(len(r), dtype=np.int64) - start[r]
    keep = rank < lengths[r]
    r, c = r[keep], c[keep]
    assert np.array_equal(np.bincount(r, minlength=len(lengths)), lengths), "over-draw too small for a row"
    order = np.argsort(r * cols + c, kind="stable")
    return r[order], c[order]


C1_SHAPE = (47894, 41550, 1790490)  # SuiteSparse Meszaros/IG5-18: rows, cols, nnz


def c1_standin(seed=1):
    """Stand-in for BASELINE.json configs[0] (matrix_example/suite_collection/IG5-18.mtx is a missing blob): the shape and
    EXACT non-zero count of IG5-18, LP-constraint-like row lengths (lognormal, mean 37.4, sigma 0.9), uniform columns
    (SURVEY.md section 8d).  Returns (rows, cols, row, col)."""
    rows, cols, nnz = C1_SHAPE
    rng = np.random.default_rng(0x6753 + seed)
    sigma = 0.9
    mu = np.log(nnz / rows) - sigma * sigma / 2
    lengths = np.maximum(1, rng.lognormal(mu, sigma, size=rows)).astype(np.int64)
    # land on the exact count: +-1 on randomly chosen rows (never below 1)
    diff = nnz - int(lengths.sum())
    while diff != 0:
        pick = rng.choice(rows, size=min(abs(diff), rows), replace=False)
        if diff < 0:
            pick = pick[lengths[pick] > 1]
        lengths[pick] += 1 if diff > 0 else -1
        diff = nnz - int(lengths.sum())
    r, c = exact_lengths_matrix(rng, lengths, cols)
    return rows, cols, r, c


def write_mtx(path, rows, cols, r, c, vals=None, comments=("%%MatrixMarket matrix coordinate real general",)):
    """1-based, row-major-sorted coordinate text, `%` comment lines first, then `rows cols nnz` -- the only input the
    reference loader accepts (struct.cc:97-125)."""
    with open(path, "w") as f:
        for line in comments:
            f.write(line + "\n")
        f.write("%d %d %d\n" % (rows, cols, len(r)))
        if vals is None:
            np.savetxt(f, np.stack([r + 1, c + 1, np.ones(len(r), dtype=np.int64)], axis=1), fmt="%d %d %d")
        else:
            rc = np.char.add(np.char.add((r + 1).astype(str), " "), (c + 1).astype(str))
            f.write("\n".join(np.char.add(np.char.add(rc, " "), np.char.mod("%.9g", vals)).tolist()) + "\n")
