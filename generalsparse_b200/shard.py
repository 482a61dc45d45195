"""Row sharding of one SpMM over G processes (one per GPU): contiguous row blocks with cut points from the reference's
balanced-interval rule (data_transform_common.cc:934-988 -> gsh_partition_rows_balanced), B replicated, C rows disjoint.
No data-path collective: a rank needs nothing from the others to compute its rows of C."""
import numpy as np

from . import api


def row_cuts(row_nnz, world):
    """G+1 begin rows; every rank computes the same cuts from the same row-length histogram"""
    return api.partition_rows_balanced(np.ascontiguousarray(row_nnz, dtype=np.uint64), world).astype(np.int64)


def shard_coo(rows, r, c, v, rank, world):
    """-> (lo, hi, r_local, c_local, v_local): the rank's row block [lo, hi) with rows rebased to 0"""
    r = np.asarray(r, dtype=np.int64)
    cuts = row_cuts(np.bincount(r, minlength=rows), world)
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    sel = (r >= lo) & (r < hi)
    return lo, hi, r[sel] - lo, np.asarray(c)[sel], (None if v is None else np.asarray(v)[sel])
