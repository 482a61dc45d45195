"""ctypes doors onto oracle/liboracle_spmm.so (the CPU restatement of the reference's SpMM) and,
where built, oracle/_ref/libref_spmm.so (the reference's own function).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing in the product path (generalsparse_b200.api / the C ABI) imports this.
"""
import ctypes as C
import os

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # repo root: this file lives in oracle/
_ORACLE = os.path.join(_ROOT, "oracle", "liboracle_spmm.so")
_REF = os.path.join(_ROOT, "oracle", "_ref", "libref_spmm.so")
_o = None
_r = None


def oracle_lib():
    global _o
    if _o is None:
        if not os.path.exists(_ORACLE):
            raise ImportError("oracle/liboracle_spmm.so is not built (make -C oracle liboracle_spmm.so)")
        _o = C.CDLL(_ORACLE)
    return _o


def ref_lib():
    """The reference's own spmm_reference_host compiled from /root/reference (None if not built)."""
    global _r
    if _r is None and os.path.exists(_REF):
        _r = C.CDLL(_REF)
    return _r


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _prep(indptr, indices, vals, B):
    return (np.ascontiguousarray(indptr, dtype=np.int64), np.ascontiguousarray(indices, dtype=np.int32),
            np.ascontiguousarray(vals, dtype=np.float32), np.ascontiguousarray(B, dtype=np.float32))


def oracle_spmm_f32(M, N, indptr, indices, vals, B, mt=False):
    indptr, indices, vals, B = _prep(indptr, indices, vals, B)
    out = np.empty((M, N), dtype=np.float32)
    fn = oracle_lib().oracle_spmm_f32_mt if mt else oracle_lib().oracle_spmm_f32
    fn(C.c_int64(M), C.c_int64(N), _p(indptr), _p(indices), _p(vals), _p(B), _p(out))
    return out


def oracle_spmm_f64acc(M, N, indptr, indices, vals, B):
    indptr, indices, vals, B = _prep(indptr, indices, vals, B)
    out = np.empty((M, N), dtype=np.float64)
    oracle_lib().oracle_spmm_f64acc(C.c_int64(M), C.c_int64(N), _p(indptr), _p(indices), _p(vals), _p(B), _p(out))
    return out


def oracle_spmm_absrow(M, N, indptr, indices, vals, B):
    indptr, indices, vals, B = _prep(indptr, indices, vals, B)
    out = np.empty((M, N), dtype=np.float64)
    oracle_lib().oracle_spmm_absrow(C.c_int64(M), C.c_int64(N), _p(indptr), _p(indices), _p(vals), _p(B), _p(out))
    return out


def ref_spmm_f32(M, N, K, indptr, indices, vals, B, threads=1):
    """The reference's own loop (arm_code/kernel_lib.hpp:859-881); threads>1 applies it to disjoint row ranges."""
    L = ref_lib()
    assert L is not None
    ip = np.ascontiguousarray(indptr, dtype=np.int32)
    ix = np.ascontiguousarray(indices, dtype=np.int32)
    v = np.ascontiguousarray(vals, dtype=np.float32)
    Bm = np.ascontiguousarray(B, dtype=np.float32)
    out = np.empty((M, N), dtype=np.float32)
    if threads == 1:
        L.ref_spmm_host_f32(M, N, K, _p(ip), _p(ix), _p(v), _p(Bm), _p(out))
    else:
        L.ref_spmm_host_f32_mt(M, N, K, _p(ip), _p(ix), _p(v), _p(Bm), _p(out), threads)
    return out


def csr_from_coo(rows, r):
    indptr = np.zeros(rows + 1, dtype=np.int64)
    np.add.at(indptr, np.asarray(r, dtype=np.int64) + 1, 1)
    return np.cumsum(indptr)
