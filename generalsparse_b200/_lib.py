"""ctypes binding of libgsparse_b200.so (include/gsparse_b200.h).

The library is built in-tree by `__graft_entry__.build()` / `make -C generalsparse_b200/csrc`.
There is NO fallback: if the shared library is missing, importing the engine fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgsparse_b200.so")

# every symbol include/gsparse_b200.h declares (tests check that the .so exports all of them)
SYMBOLS = [
    "gs_last_error", "gs_version", "gs_family_name",
    "gsh_metadata_from_mtx", "gsh_metadata_from_coo", "gsh_metadata_free", "gsh_metadata_count", "gsh_metadata_key",
    "gsh_metadata_info", "gsh_metadata_read_u64", "gsh_metadata_read_f64", "gsh_metadata_set_vals", "gsh_metadata_dump",
    "gsh_metadata_check", "gsh_config_load", "gsh_config_set", "gsh_config_get",
    "gsh_session_create", "gsh_session_free", "gsh_session_add_and_run", "gsh_session_record_count", "gsh_session_record",
    "gsh_session_run_family", "gsh_session_compile", "gsh_session_plan_desc", "gsh_partition_rows_balanced",
    "gsh_tile_stream_roundtrip", "gsh_sp24_roundtrip", "gsh_sp24_tiled_roundtrip",
    "gs_create", "gs_destroy", "gs_stream", "gs_upload_array", "gs_upload_session", "gs_upload_session_as", "gs_plan_create", "gs_plan_destroy",
    "gs_plan_set_kernel", "gs_set_dense", "gs_set_dense_device", "gs_launch", "gs_run", "gs_get_output", "gs_output_device",
    "gs_spmm_host", "gs_spmm_host_async", "gs_synchronize", "gs_plan_algorithmic_bytes", "gs_plan_launch_count", "gs_plan_kernel_name", "gs_plan_sp24_probe", "gs_execute",
    "gs_launch_no_alloc", "gs_plan_format_bytes", "gsh_metadata_row_block", "gsh_metadata_row_nnz", "gs_plan_set_output_device",
    "gs_comm_unique_id", "gs_comm_init", "gs_comm_allgather_rows", "gs_comm_destroy",
    "gs_exec_create", "gs_exec_destroy", "gs_exec_devices", "gs_exec_build", "gs_exec_row_cuts", "gs_exec_kernel_name", "gs_exec_set_dense", "gs_exec_run",
    "gsh_session_generate_final_program", "gs_execute_binary", "gsh_metadata_fit_model", "gs_plan_model_bytes",
    "gs_transform_device", "gs_transform_host_coo", "gs_download_array", "gs_exec_build_on_device", "gs_transform_device_merge_path",
    "gs_exec_get_output", "gs_exec_enable_allgather", "gs_exec_allgather_output", "gs_exec_get_replica",
]

(GS_U8, GS_U16, GS_U32, GS_U64, GS_F32, GS_F64, GS_F16) = range(7)
(GLOBAL_META, TBLOCK_META, WARP_META, THREAD_META) = range(4)
FAMILIES = ["thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment", "block_total", "block_bitmap",
            "b200_row_smem", "b200_tcgen05", "b200_nnz", "b200_sp24", "b200_merge"]
FAMILY_ID = {n: i for i, n in enumerate(FAMILIES)}
FAMILY_ID["auto"] = -1  # GS_KERNEL_AUTO
FAMILY_ID["chain"] = -2  # GS_KERNEL_CHAIN
STATUS = {0: "GS_OK", 1: "GS_ERR_INVALID", 2: "GS_ERR_CUDA", 3: "GS_ERR_UNSUPPORTED", 4: "GS_ERR_NOT_FOUND",
          5: "GS_ERR_PADDING_RATE", 6: "GS_ERR_INTERNAL"}


class PlanDesc(C.Structure):
    _fields_ = [("family", C.c_int32), ("sub_matrix_id", C.c_int32), ("sparse_coarsen", C.c_int32), ("coarsen", C.c_int32),
                ("vector_width", C.c_int32), ("need_warp_reduction", C.c_int32), ("thread_for_row", C.c_int32),
                ("has_row_permutation", C.c_int32), ("grid", C.c_uint32 * 2), ("block", C.c_uint32 * 2),
                ("rows", C.c_uint64), ("cols", C.c_uint64), ("origin_nnz", C.c_uint64), ("stored_nnz", C.c_uint64),
                ("n_deps", C.c_int32), ("deps", (C.c_char * 64) * 16)]


class GsError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("%s: %s" % (STATUS.get(status, status), msg))
        self.status = status


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libgsparse_b200.so is not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "-- there is no CPU or PyTorch fallback for the SpMM path" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, cp, i32, u64, sz = C.c_void_p, C.c_char_p, C.c_int, C.c_uint64, C.c_size_t
    P = C.POINTER
    sig = {
        "gs_last_error": (cp, []), "gs_version": (cp, []), "gs_family_name": (cp, [i32]),
        "gsh_metadata_from_mtx": (i32, [cp, cp, i32, P(vp)]),
        "gsh_metadata_from_coo": (i32, [u64, u64, u64, vp, vp, vp, P(vp)]),
        "gsh_metadata_free": (None, [vp]), "gsh_metadata_count": (u64, [vp]),
        "gsh_metadata_key": (i32, [vp, u64, cp, sz]),
        "gsh_metadata_info": (i32, [vp, cp, P(u64), P(i32), P(i32)]),
        "gsh_metadata_read_u64": (i32, [vp, cp, vp, u64]), "gsh_metadata_read_f64": (i32, [vp, cp, vp, u64]),
        "gsh_metadata_set_vals": (i32, [vp, i32, vp, u64]), "gsh_metadata_dump": (i32, [vp, cp]),
        "gsh_metadata_check": (i32, [vp]),
        "gsh_config_load": (i32, [cp]), "gsh_config_set": (i32, [cp, cp]), "gsh_config_get": (i32, [cp, cp, sz]),
        "gsh_session_create": (i32, [vp, i32, P(vp)]), "gsh_session_free": (None, [vp]),
        "gsh_session_add_and_run": (i32, [vp, cp, vp, i32]), "gsh_session_record_count": (u64, [vp]),
        "gsh_session_record": (i32, [vp, u64, cp, sz]), "gsh_session_run_family": (i32, [vp, i32, i32, i32, i32]),
        "gsh_session_compile": (i32, [vp]), "gsh_session_plan_desc": (i32, [vp, P(PlanDesc)]),
        "gsh_partition_rows_balanced": (i32, [vp, u64, i32, vp]),
        "gsh_tile_stream_roundtrip": (i32, [vp, vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp]),
        "gs_create": (i32, [P(vp), i32]), "gs_destroy": (None, [vp]), "gs_stream": (vp, [vp]),
        "gs_upload_array": (i32, [vp, cp, vp, u64, i32, i32]), "gs_upload_session": (i32, [vp, vp, i32]), "gs_upload_session_as": (i32, [vp, vp, i32, i32]),
        "gs_plan_create": (i32, [vp, P(PlanDesc), i32, i32, P(vp)]), "gs_plan_destroy": (None, [vp]),
        "gs_transform_device": (i32, [vp, i32, u64, u64, u64, vp, vp, vp, i32, i32, i32, i32, i32, P(PlanDesc)]),
        "gs_transform_host_coo": (i32, [vp, i32, u64, u64, u64, vp, vp, vp, i32, i32, i32, i32, i32, P(PlanDesc), P(C.c_float)]),
        "gs_download_array": (i32, [vp, cp, vp, u64, P(u64), P(i32)]),
        "gs_transform_device_merge_path": (i32, [vp, i32, C.c_uint32, i32, u64, u64, u64, vp, vp, vp, i32, i32, i32, i32, vp, vp, i32, i32, P(PlanDesc)]),
        "gs_plan_set_kernel": (i32, [vp, i32]), "gs_set_dense": (i32, [vp, vp, i32]), "gs_set_dense_device": (i32, [vp, vp]),
        "gs_launch": (i32, [vp]), "gs_run": (i32, [vp, i32, P(C.c_float)]), "gs_get_output": (i32, [vp, vp]),
        "gs_output_device": (vp, [vp]), "gs_spmm_host": (i32, [vp, vp, vp]),
        "gs_spmm_host_async": (i32, [vp, vp, vp]), "gs_synchronize": (i32, [vp]),
        "gs_plan_algorithmic_bytes": (u64, [vp]), "gs_plan_format_bytes": (u64, [vp]), "gs_plan_launch_count": (i32, [vp]), "gs_plan_kernel_name": (cp, [vp]),
        "gs_execute": (i32, [vp, i32, P(C.c_float), P(C.c_float)]),
        "gsh_sp24_roundtrip": (i32, [vp, vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp]),
        "gsh_sp24_tiled_roundtrip": (i32, [vp, vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp]),
        "gs_plan_sp24_probe": (i32, [vp, vp]),
        "gs_launch_no_alloc": (i32, [vp, vp, vp, vp]),
        "gsh_metadata_row_block": (i32, [vp, u64, u64, P(vp)]), "gsh_metadata_row_nnz": (i32, [vp, vp, u64]),
        "gs_plan_set_output_device": (i32, [vp, vp]),
        "gsh_session_generate_final_program": (i32, [vp, C.c_uint, P(u64), cp, sz]),
        "gsh_metadata_fit_model": (i32, [vp, cp, P(i32), vp]), "gs_plan_model_bytes": (u64, [vp]),
        "gs_execute_binary": (i32, [cp, cp, i32, i32, P(C.c_float), P(C.c_float)]),
        "gs_comm_unique_id": (i32, [vp]), "gs_comm_init": (i32, [vp, i32, i32, vp]), "gs_comm_allgather_rows": (i32, [vp, vp, vp, u64]),
        "gs_comm_destroy": (None, [vp]),
        "gs_exec_create": (i32, [P(vp), vp, i32]), "gs_exec_destroy": (None, [vp]), "gs_exec_devices": (i32, [vp]),
        "gs_exec_build": (i32, [vp, vp, i32, i32, i32, i32, i32, i32]), "gs_exec_build_on_device": (i32, [vp, vp, i32, i32, i32, i32, i32, i32]), "gs_exec_row_cuts": (i32, [vp, vp, vp]), "gs_exec_kernel_name": (cp, [vp, i32]),
        "gs_exec_set_dense": (i32, [vp, vp, i32]), "gs_exec_run": (i32, [vp, i32, P(C.c_float), vp]), "gs_exec_get_output": (i32, [vp, vp]),
        "gs_exec_enable_allgather": (i32, [vp]), "gs_exec_allgather_output": (i32, [vp, P(C.c_float)]), "gs_exec_get_replica": (i32, [vp, i32, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(status):
    if status != 0:
        raise GsError(status, lib().gs_last_error().decode())
