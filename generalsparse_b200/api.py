"""Python mirror of the reference's operator-facing interface for the SpMM path.

Thin wrappers over the C ABI (include/gsparse_b200.h): the work happens in libgsparse_b200.so
(C++ operator graph + CUDA kernels).  Names follow the reference: meta_data_set
(metadata_set.hpp:103), code_generator/operator_executer (code_generator.hpp:48,
operator_executer.hpp:10), test_spmm_<family> chains (token_test.cc:1003-1582),
execute_binary (executor.hpp:15).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import FAMILIES, FAMILY_ID, GsError, PlanDesc, check, lib

_NP2GS = {np.dtype(np.uint8): _lib.GS_U8, np.dtype(np.uint16): _lib.GS_U16, np.dtype(np.uint32): _lib.GS_U32,
          np.dtype(np.uint64): _lib.GS_U64, np.dtype(np.float32): _lib.GS_F32, np.dtype(np.float64): _lib.GS_F64,
          np.dtype(np.float16): _lib.GS_F16}
_GS2NP = {v: k for k, v in _NP2GS.items()}
_GS2NAME = {0: "UNSIGNED_CHAR", 1: "UNSIGNED_SHORT", 2: "UNSIGNED_INT", 3: "UNSIGNED_LONG", 4: "FLOAT", 5: "DOUBLE"}


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def config_set(key, value):
    if isinstance(value, bool):
        v = "true" if value else "false"
    elif isinstance(value, (int, float)):
        v = repr(value)
    else:
        v = '"%s"' % value
    check(lib().gsh_config_set(key.encode(), v.encode()))


def config_get(key):
    buf = C.create_string_buffer(512)
    check(lib().gsh_config_get(key.encode(), buf, 512))
    return buf.value.decode()


class MetaDataSet:
    """meta_data_set of one matrix (create_init_metadata_set_from_file, metadata_set.cc:612)."""

    def __init__(self, handle):
        self._h = handle

    @classmethod
    def from_mtx(cls, path, float_precise="float", reference_ones=True):
        h = C.c_void_p()
        check(lib().gsh_metadata_from_mtx(path.encode(), float_precise.encode(), int(reference_ones), C.byref(h)))
        return cls(h)

    @classmethod
    def from_coo(cls, rows, cols, row, col, vals=None):
        row = np.ascontiguousarray(row, dtype=np.int64)
        col = np.ascontiguousarray(col, dtype=np.int64)
        v = None if vals is None else np.ascontiguousarray(vals, dtype=np.float32)
        h = C.c_void_p()
        check(lib().gsh_metadata_from_coo(rows, cols, len(row), _ptr(row), _ptr(col), None if v is None else _ptr(v), C.byref(h)))
        return cls(h)

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # (module globals are gone at interpreter shutdown: the process takes the handle with it)
            lib().gsh_metadata_free(self._h)
            self._h = None

    def keys(self):
        n = lib().gsh_metadata_count(self._h)
        buf = C.create_string_buffer(256)
        out = []
        for i in range(n):
            check(lib().gsh_metadata_key(self._h, i, buf, 256))
            out.append(buf.value.decode())
        return out

    def info(self, key):
        n, m, c = C.c_uint64(), C.c_int(), C.c_int()
        check(lib().gsh_metadata_info(self._h, key.encode(), C.byref(n), C.byref(m), C.byref(c)))
        return n.value, _GS2NAME[m.value], _GS2NAME[c.value]

    def __contains__(self, key):
        return key in self.keys()

    def __getitem__(self, key):
        n, mem, _ = self.info(key)
        if mem in ("FLOAT", "DOUBLE"):
            a = np.empty(n, dtype=np.float64)
            check(lib().gsh_metadata_read_f64(self._h, key.encode(), _ptr(a), n))
        else:
            a = np.empty(n, dtype=np.uint64)
            check(lib().gsh_metadata_read_u64(self._h, key.encode(), _ptr(a), n))
        return a

    def fit_model(self, key):
        """model-driven index compression: (type name, coefficient, intercept, cycle or branch count) of the closed form the item follows"""
        t, prm = C.c_int(), np.zeros(4, dtype=np.uint64)
        check(lib().gsh_metadata_fit_model(self._h, key.encode(), C.byref(t), _ptr(prm)))
        return ["none", "linear", "constant", "branch", "cycle_linear", "cycle_increase"][t.value], int(prm[0]), int(prm[1]), int(prm[2])

    def set_vals(self, vals, sub_matrix_id=0):
        v = np.ascontiguousarray(vals, dtype=np.float32)
        check(lib().gsh_metadata_set_vals(self._h, sub_matrix_id, _ptr(v), len(v)))

    def row_block(self, lo, hi):
        """rows [lo, hi) of the matrix as loaded, rebased to 0: one GPU's block of the row partition"""
        h = C.c_void_p()
        check(lib().gsh_metadata_row_block(self._h, lo, hi, C.byref(h)))
        return MetaDataSet(h)

    def row_nnz(self):
        rows = int(self["GLOBAL_META_origin_row_num_-1"][0])
        out = np.empty(rows, dtype=np.uint64)
        check(lib().gsh_metadata_row_nnz(self._h, _ptr(out), rows))
        return out

    def dump(self, directory):
        check(lib().gsh_metadata_dump(self._h, directory.encode()))

    def check(self):
        check(lib().gsh_metadata_check(self._h))


class Session:
    """code_generator + operator_executer of one sub-matrix."""

    def __init__(self, ms, sub_matrix_id=0):
        self.ms = ms
        h = C.c_void_p()
        check(lib().gsh_session_create(ms._h, sub_matrix_id, C.byref(h)))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # (module globals are gone at interpreter shutdown: the process takes the handle with it)
            lib().gsh_session_free(self._h)
            self._h = None

    def add_and_run(self, op_name, *args):
        a = np.array([int(x) for x in args], dtype=np.int64)
        check(lib().gsh_session_add_and_run(self._h, op_name.encode(), _ptr(a) if len(a) else None, len(a)))
        return self

    def run_family(self, family, N, sparse_coarsen=1, coarsen=1):
        fid = FAMILY_ID[family] if isinstance(family, str) else family
        check(lib().gsh_session_run_family(self._h, fid, N, sparse_coarsen, coarsen))
        return self

    def compile(self):
        check(lib().gsh_session_compile(self._h))
        return self

    def records(self):
        n = lib().gsh_session_record_count(self._h)
        buf = C.create_string_buffer(1024)
        out = []
        for i in range(n):
            check(lib().gsh_session_record(self._h, i, buf, 1024))
            out.append(buf.value.decode())
        return out

    def generate_final_program(self, repeat=10000):
        """code_generator::generate_final_program: -> (id, directory) of the hand-off (one text file per array the kernel reads)"""
        i, buf = C.c_uint64(), C.create_string_buffer(1024)
        check(lib().gsh_session_generate_final_program(self._h, repeat, C.byref(i), buf, 1024))
        return i.value, buf.value.decode()

    def plan_desc(self):
        d = PlanDesc()
        check(lib().gsh_session_plan_desc(self._h, C.byref(d)))
        return d

    def deps(self):
        d = self.plan_desc()
        return [bytes(d.deps[i]).split(b"\0")[0].decode() for i in range(d.n_deps)]


def execute_binary(execute_path, matrix_file, repeat=10000, kernel="auto"):
    """execute_binary(path, matrix, &ms_total, &gflops) of executor.hpp:15 on a hand-off directory -> (ms_total, gflops)"""
    ms, gf = C.c_float(), C.c_float()
    check(lib().gs_execute_binary(execute_path.encode(), matrix_file.encode(), repeat, FAMILY_ID[kernel] if isinstance(kernel, str) else kernel, C.byref(ms), C.byref(gf)))
    return ms.value, gf.value


def partition_rows_balanced(row_nnz, G):
    rn = np.ascontiguousarray(row_nnz, dtype=np.uint64)
    out = np.empty(G + 1, dtype=np.uint64)
    check(lib().gsh_partition_rows_balanced(_ptr(rn), len(rn), G, _ptr(out)))
    return out


class Engine:
    """Device context (gs_ctx): replaces executor.cc's compile-and-fork."""

    def __init__(self, device_id=0):
        h = C.c_void_p()
        check(lib().gs_create(C.byref(h), device_id))
        self._h = h
        self.device_id = device_id

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # (module globals are gone at interpreter shutdown: the process takes the handle with it)
            lib().gs_destroy(self._h)
            self._h = None

    @property
    def stream(self):
        return lib().gs_stream(self._h)

    def synchronize(self):
        check(lib().gs_synchronize(self._h))

    # ---- optional all-gather of C over NCCL, one process per GPU ----
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        check(lib().gs_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, nranks, rank, unique_id):
        check(lib().gs_comm_init(self._h, nranks, rank, C.c_char_p(unique_id)))

    def comm_allgather_rows(self, c_full_ptr, row_cuts, row_bytes):
        cuts = np.ascontiguousarray(row_cuts, dtype=np.uint64)
        check(lib().gs_comm_allgather_rows(self._h, C.c_void_p(c_full_ptr), _ptr(cuts), row_bytes))

    def upload_array(self, key, arr, half=False):
        a = np.ascontiguousarray(arr)
        check(lib().gs_upload_array(self._h, key.encode(), _ptr(a), a.size, _NP2GS[a.dtype], int(half)))

    def upload_session(self, session, half=False, as_id=-1):
        check(lib().gs_upload_session_as(self._h, session._h, int(half), as_id))

    def transform_coo(self, family, rows, cols, r, c, vals, N, sparse_coarsen=1, coarsen=1, half=False, as_id=0):
        """ the driver chain of `family` ("thread_total", "warp_total", "thread_bit_map") built by device passes from a row-major
        COO; returns (PlanDesc, device milliseconds of the passes).  The arrays land where upload_session(as_id=...) would have put them."""
        r32, c32 = np.ascontiguousarray(r, dtype=np.uint32), np.ascontiguousarray(c, dtype=np.uint32)
        v32 = np.ones(len(r32), np.float32) if vals is None else np.ascontiguousarray(vals, dtype=np.float32)
        desc, ms = PlanDesc(), C.c_float(0)
        fid = FAMILY_ID[family] if isinstance(family, str) else family
        check(lib().gs_transform_host_coo(self._h, fid, rows, cols, len(r32), _ptr(r32), _ptr(c32), _ptr(v32), N, sparse_coarsen, coarsen, int(half), as_id, C.byref(desc),
                                          C.byref(ms)))
        return desc, ms.value

    def transform_coo_device(self, family, rows, cols, nnz, row_ptr, col_ptr, val_ptr, N, sparse_coarsen=1, coarsen=1, half=False, as_id=0):
        """the same from device pointers (u32 row ids ascending, u32 column ids, fp32 values)"""
        desc = PlanDesc()
        fid = FAMILY_ID[family] if isinstance(family, str) else family
        check(lib().gs_transform_device(self._h, fid, rows, cols, nnz, C.c_void_p(row_ptr), C.c_void_p(col_ptr), C.c_void_p(val_ptr), N, sparse_coarsen, coarsen, int(half), as_id,
                                        C.byref(desc)))
        return desc

    def transform_coo_merge_path(self, level, work_size, rows, cols, r, c, vals, N, coarsen=1, vector_width=None, grid=(64, 1), block=(8, 4), sort=False, half=False, as_id=0):
        """[sort_operator +] merge_path_{thread,warp,tblock}_operator(work_size) + the level's total reduction, built by device passes; level: 0 / 1 / 2
        or "thread" / "warp" / "tblock"; grid / block / vector_width: the launch-shape hints of the chain's grid_block_operator"""
        lv = {"thread": 0, "warp": 1, "tblock": 2}.get(level, level)
        r32, c32 = np.ascontiguousarray(r, dtype=np.uint32), np.ascontiguousarray(c, dtype=np.uint32)
        v32 = np.ones(len(r32), np.float32) if vals is None else np.ascontiguousarray(vals, dtype=np.float32)
        g, b = np.asarray(grid, np.uint32), np.asarray(block, np.uint32)
        desc = PlanDesc()
        check(lib().gs_transform_device_merge_path(self._h, lv, work_size, int(sort), rows, cols, len(r32), _ptr(r32), _ptr(c32), _ptr(v32), 1, N, coarsen,
                                                   min(N, 32) if vector_width is None else vector_width, _ptr(g), _ptr(b), int(half), as_id, C.byref(desc)))
        return desc

    def download_array(self, key):
        """a device array of the context as numpy, in its device dtype"""
        n, dt = C.c_uint64(0), C.c_int32(0)
        check(lib().gs_download_array(self._h, key.encode(), None, 0, C.byref(n), C.byref(dt)))
        out = np.empty(n.value, dtype=_GS2NP[dt.value])
        check(lib().gs_download_array(self._h, key.encode(), _ptr(out), out.nbytes, C.byref(n), C.byref(dt)))
        return out

    def plan(self, session_or_desc, N, half=False, kernel=None, sub_matrix_id=None):
        desc = session_or_desc.plan_desc() if isinstance(session_or_desc, Session) else session_or_desc
        if sub_matrix_id is not None:
            desc.sub_matrix_id = sub_matrix_id
        h = C.c_void_p()
        check(lib().gs_plan_create(self._h, C.byref(desc), N, int(half), C.byref(h)))
        p = Plan(self, h, desc, N, half)
        if kernel is not None:
            p.set_kernel(kernel)
        return p


class Plan:
    def __init__(self, engine, handle, desc, N, half):
        self.engine, self._h, self.desc, self.N, self.half = engine, handle, desc, N, half
        self.vdtype = np.float16 if half else np.float32

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # (module globals are gone at interpreter shutdown: the process takes the handle with it)
            lib().gs_plan_destroy(self._h)
            self._h = None

    def set_kernel(self, family):
        fid = FAMILY_ID[family] if isinstance(family, str) else family
        check(lib().gs_plan_set_kernel(self._h, fid))

    def set_dense(self, B):
        B = np.ascontiguousarray(B)
        assert B.shape == (self.desc.cols, self.N), (B.shape, self.desc.cols, self.N)
        check(lib().gs_set_dense(self._h, _ptr(B), _NP2GS[B.dtype]))

    def set_dense_device(self, ptr):
        check(lib().gs_set_dense_device(self._h, C.c_void_p(ptr)))

    def launch(self):
        check(lib().gs_launch(self._h))

    def run(self, repeat=1):
        ms = C.c_float()
        check(lib().gs_run(self._h, repeat, C.byref(ms)))
        return ms.value

    def output(self):
        out = np.empty((self.desc.rows, self.N), dtype=self.vdtype)
        check(lib().gs_get_output(self._h, _ptr(out)))
        return out

    def set_output_device(self, ptr):
        check(lib().gs_plan_set_output_device(self._h, C.c_void_p(ptr)))

    def launch_no_alloc(self, b_ptr, c_ptr, stream_ptr=None):
        """gs_launch_no_alloc: caller's device B, C and stream; scratch in the library's static device array"""
        check(lib().gs_launch_no_alloc(self._h, C.c_void_p(b_ptr), C.c_void_p(c_ptr), C.c_void_p(stream_ptr)))

    def output_device(self):
        return lib().gs_output_device(self._h)

    def spmm_host(self, B, out=None):
        B = np.ascontiguousarray(B, dtype=self.vdtype)
        if out is None:
            out = np.empty((self.desc.rows, self.N), dtype=self.vdtype)
        check(lib().gs_spmm_host(self._h, _ptr(B), _ptr(out)))
        return out

    def spmm_host_ptr(self, b_ptr, c_ptr):
        check(lib().gs_spmm_host(self._h, C.c_void_p(b_ptr), C.c_void_p(c_ptr)))

    def spmm_host_async_ptr(self, b_ptr, c_ptr):
        """pinned host pointers; C is complete after Engine.synchronize()"""
        check(lib().gs_spmm_host_async(self._h, C.c_void_p(b_ptr), C.c_void_p(c_ptr)))

    def execute(self, repeat):
        """execute_binary(path, matrix, &ms_total, &gflops) of executor.hpp:15"""
        ms, gf = C.c_float(), C.c_float()
        check(lib().gs_execute(self._h, repeat, C.byref(ms), C.byref(gf)))
        return ms.value, gf.value

    @property
    def algorithmic_bytes(self):
        return lib().gs_plan_algorithmic_bytes(self._h)

    @property
    def model_bytes(self):
        return lib().gs_plan_model_bytes(self._h)

    @property
    def format_bytes(self):
        return lib().gs_plan_format_bytes(self._h)

    @property
    def launch_count(self):
        return lib().gs_plan_launch_count(self._h)

    def sp24_probe(self):
        """the 4 metadata columns of the first tile as tensor memory holds them (plan created under GS_SP24_PROBE=1)"""
        out = np.empty((128, 4), dtype=np.uint32)
        check(lib().gs_plan_sp24_probe(self._h, _ptr(out)))
        return out

    @property
    def kernel_name(self):
        return lib().gs_plan_kernel_name(self._h).decode()


class Executor:
    """gs_exec: one SpMM over several GPUs from one process (north star (3)): nnz-balanced row blocks, B replicated, C rows disjoint"""

    def __init__(self, device_ids):
        ids = np.ascontiguousarray(device_ids, dtype=np.int32)
        h = C.c_void_p()
        check(lib().gs_exec_create(C.byref(h), _ptr(ids), len(ids)))
        self._h, self.n = h, len(ids)

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # (module globals are gone at interpreter shutdown: the process takes the handle with it)
            lib().gs_exec_destroy(self._h)
            self._h = None

    def build(self, ms, family, N, sparse_coarsen=1, coarsen=1, half=False, kernel="auto", on_device=False):
        """on_device: every block's chain is built by its GPU (gs_exec_build_on_device) instead of the host pipeline"""
        fid = FAMILY_ID[family] if isinstance(family, str) else family
        kid = FAMILY_ID[kernel] if isinstance(kernel, str) else kernel
        check((lib().gs_exec_build_on_device if on_device else lib().gs_exec_build)(self._h, ms._h, fid, N, sparse_coarsen, coarsen, int(half), kid))
        self.N, self.half = N, half
        self.rows = int(ms["GLOBAL_META_origin_row_num_-1"][0])
        self.cols = int(ms["GLOBAL_META_origin_col_num_-1"][0])
        return self

    def row_cuts(self):
        cuts, nnz = np.empty(self.n + 1, dtype=np.uint64), np.empty(self.n, dtype=np.uint64)
        check(lib().gs_exec_row_cuts(self._h, _ptr(cuts), _ptr(nnz)))
        return cuts, nnz

    def kernel_name(self, g=0):
        return lib().gs_exec_kernel_name(self._h, g).decode()

    def set_dense(self, B):
        B = np.ascontiguousarray(B)
        assert B.shape == (self.cols, self.N)
        check(lib().gs_exec_set_dense(self._h, _ptr(B), _NP2GS[B.dtype]))

    def run(self, repeat=1):
        mx, per = C.c_float(), np.zeros(self.n, dtype=np.float32)
        check(lib().gs_exec_run(self._h, repeat, C.byref(mx), _ptr(per)))
        return mx.value, per

    def output(self):
        out = np.empty((self.rows, self.N), dtype=np.float16 if self.half else np.float32)
        check(lib().gs_exec_get_output(self._h, _ptr(out)))
        return out

    def enable_allgather(self):
        check(lib().gs_exec_enable_allgather(self._h))

    def allgather_output(self):
        ms = C.c_float()
        check(lib().gs_exec_allgather_output(self._h, C.byref(ms)))
        return ms.value

    def replica(self, g):
        out = np.empty((self.rows, self.N), dtype=np.float16 if self.half else np.float32)
        check(lib().gs_exec_get_replica(self._h, g, _ptr(out)))
        return out


__all__ = ["Executor", "MetaDataSet", "Session", "Engine", "Plan", "GsError", "FAMILIES", "FAMILY_ID", "config_set", "config_get",
           "partition_rows_balanced"]
