/* gsparse_b200.h -- C ABI of libgsparse_b200.so, the B200-native SpMM engine behind
 * GeneralSparse's operator API.
 *
 * The reference has no FFI: its boundary is C++ objects + text files + a child process
 * (SURVEY.md section 8b).  This header is the C-ABI that replaces that hand-off; every entry
 * point cites the reference interface it stands in for (paths relative to the reference root).
 *
 *   gsh_*  host side: metadata store, operator graph, transform steps (no GPU needed)
 *   gs_*   device side: array upload, kernel plans, launches, timing (needs a B200)
 *
 * All functions return gs_status (0 = ok) unless stated; on failure gs_last_error() holds the
 * message.  Nothing aborts the process (the reference uses assert(false)/exit(-1),
 * executor.cc:16-19,85).  Handles are not thread-safe; use one context per host thread.
 */
#ifndef GSPARSE_B200_H
#define GSPARSE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gs_status {
    GS_OK = 0,
    GS_ERR_INVALID = 1,      /* bad argument / violated operator pre-condition (reference: assert) */
    GS_ERR_CUDA = 2,         /* CUDA runtime failure, or no usable device */
    GS_ERR_UNSUPPORTED = 3,  /* valid in the reference API, not built here */
    GS_ERR_NOT_FOUND = 4,    /* metadata key / handle missing */
    GS_ERR_PADDING_RATE = 5, /* PADDING_RATE_UP_BOUND exceeded (modify_*_by_*_pad*.cc abort) */
    GS_ERR_INTERNAL = 6
} gs_status;

typedef enum gs_dtype { GS_U8 = 0, GS_U16 = 1, GS_U32 = 2, GS_U64 = 3, GS_F32 = 4, GS_F64 = 5, GS_F16 = 6 } gs_dtype;

/* POS_TYPE (metadata_set.hpp:12-22) */
typedef enum gs_pos { GS_GLOBAL_META = 0, GS_TBLOCK_META = 1, GS_WARP_META = 2, GS_THREAD_META = 3 } gs_pos;

/* The seven kernel families the reference driver wires (token_test.cc:1003-1582), i.e. the
 * reduction tokens of reduction_token.hpp:7-110, plus the B200-native row kernels that consume
 * the same BMT/BMW arrays. */
typedef enum gs_family {
    GS_FAMILY_THREAD_TOTAL = 0,  /* total_BMT_result_reduce_to_one_register_token.cc:470-913 */
    GS_FAMILY_WARP_TOTAL = 1,    /* total_warp_result_reduce_to_one_register_token.cc:525-840 */
    GS_FAMILY_WARP_BITMAP = 2,   /* warp_bit_map_reduce_token.cc:94-392 */
    GS_FAMILY_THREAD_BITMAP = 3, /* thread_bit_map_reduce_to_two_register_token.cc:819-1616 */
    GS_FAMILY_WARP_SEGMENT = 4,  /* warp_segment_reduce_token.cc:94-440 */
    GS_FAMILY_TBLOCK_TOTAL = 5,  /* total_block_reduce_to_one_register_token.cc:374-525 */
    GS_FAMILY_TBLOCK_BITMAP = 6, /* tblock_bit_map_reduce_token.cc:83-189 */
    GS_FAMILY_B200_ROW_SMEM = 7, /* persistent row kernel, dense B panel staged in shared memory by bulk copy */
    GS_FAMILY_B200_TCGEN05 = 8,  /* sparse-tile decompress to shared memory + tcgen05.mma with TMEM accumulators */
    GS_FAMILY_B200_NNZ = 9,      /* warp-per-run kernel for the nnz-direction bit-map chains (thread_bit_map / warp_segment), N % 32 == 0 */
    GS_FAMILY_B200_SP24 = 10,    /* 2:4 sparse tensor cores: tcgen05.mma.sp over a pre-laid-out image, what the pattern cannot hold goes to a gather pass */
    GS_FAMILY_B200_MERGE = 11,   /* range kernel: merge-path parts, nnz-balanced row groups, column-direction chunks, row blocks > 1 row; N in {1,2,4,8,16} or 32k <= 512 */
    GS_FAMILY_COUNT = 12
} gs_family;
/* gs_plan_set_kernel(plan, GS_KERNEL_AUTO): the library picks the kernel for (chain, matrix, N, value type) from a cost model of the
 * measured rates -- the slice of the reference's unbuilt calculation_method_choose_operator (README.md:36, "under development") this
 * path needs; the chain's own family is the last resort */
#define GS_KERNEL_AUTO (-1)
#define GS_KERNEL_CHAIN (-2) /* gs_exec_build: keep the chain's own reference-shaped family */

const char *gs_last_error(void);
const char *gs_version(void);
const char *gs_family_name(int family);

/* ------------------------------------------------------------------ host: metadata store ---- */
typedef struct gsh_metadata gsh_metadata; /* meta_data_set, metadata_set.hpp:103-153 */

/* create_init_metadata_set_from_file, metadata_set.cc:612-707 (+ loader struct.cc:49-261).
 * reference_ones != 0 forces every value to 1 as the reference does (struct.cc:188-195). */
gs_status gsh_metadata_from_mtx(const char *path, const char *float_precise, int reference_ones, gsh_metadata **out);
/* the same store from in-memory COO (row-major sorted, 0-based); vals may be NULL (all ones) */
gs_status gsh_metadata_from_coo(uint64_t rows, uint64_t cols, uint64_t nnz, const int64_t *row, const int64_t *col,
                                const float *vals, gsh_metadata **out);
/* rows [lo, hi) of the matrix as loaded (sub-matrix 0, before any operator ran) as a store of its own, rows rebased to 0: the
 * block one GPU owns in the row partition (the reference's matrix-div operators cut sub-matrices the same way,
 * fixed_div_row_indices.cc).  gsh_metadata_row_nnz: non-zeros per row of the matrix as loaded (len rows = origin_row_num). */
gs_status gsh_metadata_row_block(const gsh_metadata *ms, uint64_t lo, uint64_t hi, gsh_metadata **out);
gs_status gsh_metadata_row_nnz(const gsh_metadata *ms, uint64_t *out, uint64_t rows);
void gsh_metadata_free(gsh_metadata *ms);
/* number of items; key of item i in insertion order ("<POS>_<name>_<id>", metadata_set.cc:147-151) */
uint64_t gsh_metadata_count(const gsh_metadata *ms);
gs_status gsh_metadata_key(const gsh_metadata *ms, uint64_t i, char *buf, size_t buf_len);
/* length, in-memory dtype (GS_U64/GS_F32/GS_F64) and compressed dtype (code_source_data.cc:383-416) */
gs_status gsh_metadata_info(const gsh_metadata *ms, const char *key, uint64_t *len, int *mem_dtype, int *compressed_dtype);
/* copy an item out as u64 (integers) or f64 (values) */
gs_status gsh_metadata_read_u64(const gsh_metadata *ms, const char *key, uint64_t *dst, uint64_t cap);
gs_status gsh_metadata_read_f64(const gsh_metadata *ms, const char *key, double *dst, uint64_t cap);
/* replace the values of sub-matrix `id` (same length) -- how real weights enter after the pattern was loaded */
gs_status gsh_metadata_set_vals(gsh_metadata *ms, int sub_matrix_id, const float *vals, uint64_t n);
/* meta_data_set::output_2_dir / meta_data_item::output_2_file (metadata_set.cc:482-515, struct.cc:1991) */
gs_status gsh_metadata_dump(const gsh_metadata *ms, const char *dir);
gs_status gsh_metadata_check(const gsh_metadata *ms); /* meta_data_set::check + logical_check */

/* Model-driven index compression (arr_optimization.hpp:13-85; the reference's MODEL_DRIVEN_INDEX_COMPRESS): the closed form an integer item
 * follows, fitted with the reference's acceptance rules (checked mode).  *type: 0 none, 1 linear (params: coefficient, intercept),
 * 2 constant (-, value), 3 branch (params[2] = number of runs), 4 cycle_linear (coefficient, intercept, cycle), 5 cycle_increase
 * (-, intercept, cycle).  params: 4 entries. */
gs_status gsh_metadata_fit_model(const gsh_metadata *ms, const char *key, int *type, uint64_t *params);

/* get_config()/set_config() (config.cc:2-40): numbers, booleans ("true"/"false") and strings */
gs_status gsh_config_load(const char *path);
gs_status gsh_config_set(const char *key, const char *json_scalar);
gs_status gsh_config_get(const char *key, char *buf, size_t buf_len);

/* ------------------------------------------------------------------ host: operator graph ---- */
/* One session = code_generator(meta_data_set, sub_matrix_id) + operator_executer
 * (code_generator.hpp:48, operator_executer.hpp:10-80). */
typedef struct gsh_session gsh_session;
gs_status gsh_session_create(gsh_metadata *ms, int sub_matrix_id, gsh_session **out);
void gsh_session_free(gsh_session *s);

/* operator_executer::add_and_run (operator_executer.cc:19-26) on the operator class named `op_name`
 * (operator.hpp; e.g. "sort_operator", "fixed_interval_row_direction_thread_blocking_operator").
 * `args` are the constructor arguments that follow code_generator_ptr in the reference's third
 * constructor flavour, in order, booleans as 0/1, POS_TYPE as gs_pos. */
gs_status gsh_session_add_and_run(gsh_session *s, const char *op_name, const int64_t *args, int n_args);
/* number of transform-step records so far and record i (convert_to_string() of the step) */
uint64_t gsh_session_record_count(const gsh_session *s);
gs_status gsh_session_record(const gsh_session *s, uint64_t i, char *buf, size_t buf_len);

/* The reference driver's operator chains, test_spmm_<family>(meta, name, sparse_coarsen, coarsen)
 * (token_test.cc:1003,1188,1253,1319,1393,1458,1518): runs the chain, sets VECTOR_WIDTH and the
 * launch shape exactly as the driver does. N = DENSE_MATRIX_SIZE. */
gs_status gsh_session_run_family(gsh_session *s, int family, int N, int sparse_coarsen, int coarsen);

/* code_generator::compile (code_generator.hpp:265): fixes the kernel family from the installed
 * reduction tokens and the list of metadata arrays the kernel reads (add_new_metadata_dependency,
 * code_generator.hpp:66). */
gs_status gsh_session_compile(gsh_session *s);

/* code_generator::generate_final_program(repeat) (code_generator.hpp:271-280): compiles if needed, writes the arrays the kernel reads to
 * ROOT_PATH_STR/data_source/<id>/ as the reference does (one text file per array, metadata_set.cc:517-571) and returns <id>; path_buf
 * (may be NULL) receives the directory.  gs_execute_binary() below runs it. */
gs_status gsh_session_generate_final_program(gsh_session *s, unsigned int repeat, uint64_t *id_out, char *path_buf, size_t path_len);

typedef struct gs_plan_desc {
    int32_t family;          /* gs_family */
    int32_t sub_matrix_id;
    int32_t sparse_coarsen;  /* (col,val) pairs per vector load; rows are padded to this multiple */
    int32_t coarsen;         /* dense columns per thread */
    int32_t vector_width;    /* VECTOR_WIDTH: threads cooperating on one row / BMTs per BMW */
    int32_t need_warp_reduction;
    int32_t thread_for_row;  /* 1: rows are disjoint -> plain stores; 0: atomicAdd (code_generator.cc:3316-3344) */
    int32_t has_row_permutation; /* original_nz_row_indices present (sort_operator ran) */
    uint32_t grid[2];        /* the reference's launch shape (grid_block_operator.cc:13-33); a hint */
    uint32_t block[2];
    uint64_t rows;           /* rows of C this sub-matrix writes = origin_row_num */
    uint64_t cols;           /* K */
    uint64_t origin_nnz;     /* unpadded nnz: the flop count uses this (code_generator.cc:614-624) */
    uint64_t stored_nnz;     /* after padding */
    int32_t n_deps;
    char deps[16][64];       /* metadata keys the kernel consumes */
} gs_plan_desc;
gs_status gsh_session_plan_desc(const gsh_session *s, gs_plan_desc *out);

/* nnz-balanced contiguous row partition for G devices: the reference's greedy rule
 * get_begin_rows_of_child_after_balance_blocking_in_row_direction (data_transform_common.cc:960-988)
 * with nnz_per_interval = ceil(nnz/G); a tail part beyond G is merged into the last.
 * row_nnz: nnz per row (len rows). out_begin_rows: G+1 entries. */
gs_status gsh_partition_rows_balanced(const uint64_t *row_nnz, uint64_t rows, int G, uint64_t *out_begin_rows);

/* B200 tile stream (csrc/host/tile_format.hpp): the device layout GS_FAMILY_B200_TCGEN05 builds from a row-direction
 * chain -- 128 x 64 tiles, each a run of (byte offset in the swizzled shared-memory image, fp16 value) pairs.  This entry
 * point builds the stream from CSR (rowptr: n_rows + 1 offsets, col, fp16 value bits), expands it again and reports
 * stats[0] = pairs stored (runs padded to 8), [1] = non-zero pairs, [2] = tiles, [3] = warp-wide 2-byte stores the
 * decompression issues, [4] = shared-memory wavefronts those stores need (bank conflicts included).
 * pass_warps: producer warps per sweep the runs are bank-ordered for (10 in the kernel; 0 = row-major runs).
 * dense_out: n_rows * n_cols fp16 bits, row-major.  Host only: the CPU tests use it, the engine does the same at plan time. */
gs_status gsh_tile_stream_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val_f16, uint32_t n_rows, uint32_t n_cols, uint32_t pass_warps,
                                    uint16_t *dense_out, uint64_t *stats);

/* B200 2:4 image (csrc/host/sp24_format.hpp): the device layout GS_FAMILY_B200_SP24 builds from a row-direction chain --
 * 128 x 128 tiles of 18 432 bytes (compressed operand in the swizzled shared-memory layout + 4-bit metadata per group of
 * four columns in tensor-memory order) plus a CSR remainder of what the 2:4 pattern cannot hold; `group` (1 or 2) row
 * panels form one unit of work and lie back to back.  This entry point
 * builds the image from CSR, expands image + remainder again and reports stats[0] = image bytes, [1] = non-zeros in
 * the image, [2] = non-zeros in the remainder, [3] = tiles.  Host only: the CPU tests use it, the engine does the
 * same at plan time. */
gs_status gsh_sp24_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val_f16, uint32_t n_rows, uint32_t n_cols, uint32_t group, uint16_t *dense_out,
                             uint64_t *stats);
/* The same with the remainder cut along the tiles (every tile image followed by the remainder entries of its 128 x 128
 * window, the form the kernel multiplies at N <= 32 without a second launch): stats[0] = bytes, [1] = non-zeros in the
 * images, [2] = non-zeros in the tile-local remainder blocks, [3] = tiles, [4] = bytes per tile, [5] = the largest entry
 * count of a tile, [6] = units left when the tiles that hold nothing leave the stream (every row panel keeps one; the
 * expansion goes through that form).  row_bytes (32, 64, 128): the bytes of one row of the 128 x N piece of B in shared memory (2 N; N = 8 runs
 * as 16), which the entries address directly. */
gs_status gsh_sp24_tiled_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val_f16, uint32_t n_rows, uint32_t n_cols, uint32_t row_bytes, uint16_t *dense_out,
                                   uint64_t *stats);

/* ------------------------------------------------------------------ device: engine ---------- */
typedef struct gs_ctx gs_ctx;   /* replaces executor.cc's compile-and-fork with an in-process device context */
typedef struct gs_plan gs_plan;

gs_status gs_create(gs_ctx **out, int device_id);
void gs_destroy(gs_ctx *ctx);
void *gs_stream(gs_ctx *ctx); /* the cudaStream_t every launch of this context goes to */

/* host array -> device (replaces the text-file hand-off metadata_set.cc:517-571 +
 * read_arr_from_file_with_data_type cuda_code/kernel_lib.hpp:704-781).  Index arrays are widened
 * to the device's canonical types (u16/u32 column ids, u32 offsets, u64 bit maps); value arrays
 * are stored as f32 or, when half != 0, as f16. */
gs_status gs_upload_array(gs_ctx *ctx, const char *key, const void *host, uint64_t len, int dtype, int half);
/* upload every array a compiled session depends on, straight from its metadata store */
gs_status gs_upload_session(gs_ctx *ctx, const gsh_session *s, int half);
/* the same, stored under sub-matrix id `as_id` (several matrices, or several copies of one, in one context);
 * a plan reaches them by setting gs_plan_desc.sub_matrix_id = as_id */
gs_status gs_upload_session_as(gs_ctx *ctx, const gsh_session *s, int half, int as_id);

/*  the chain's transform steps run on the DEVICE instead of the host pipeline: from a row-major COO already
 * on the device (u32 ids, fp32 values; row ids ascending) to the arrays a plan of `family` binds, stored under sub-matrix id `as_id`
 * exactly as gs_upload_session_as would have stored the host pipeline's -- byte for byte the same device arrays (tests/
 * test_device_transform_gpu.py).  Families: GS_FAMILY_THREAD_TOTAL = sort_operator (sort_operator.cc:70-108: get_row_order_by_length,
 * reorder_{val,col,row}_by_index, remove_empty_row_in_end_of_sub_matrix) + fixed_interval_row_direction_thread_blocking_operator(1,
 * col pad to sparse_coarsen; ...thread_blocking_operator.cc:482-565); GS_FAMILY_WARP_TOTAL / GS_FAMILY_TBLOCK_TOTAL = fixed_interval_row_
 * direction_{warp,tblock}_blocking_operator(1), no sort, no pad; GS_FAMILY_THREAD_BITMAP = fixed_interval_nnz_direction_thread_blocking_operator(32, pad) +
 * thread_bit_map_operator (thread_bit_map.cc, segment_{ptr,empty_flag,empty_row_indices}.cc); GS_FAMILY_WARP_SEGMENT = the same blocking +
 * thread_bit_map_operator at WARP level (a segment start forced at every VECTOR_WIDTH-th BMT) + warp_segment_reduce_operator (segment_offset.cc,
 * get_begin_BMTs_after_merge_thread.cc).  desc_out is what
 * gsh_session_plan_desc would return after the driver's chain (token_test.cc:1003-1092, :1188-1249, :1319-1391, :1393-1455), launch shape
 * included.  GS_FAMILY_WARP_BITMAP / GS_FAMILY_TBLOCK_BITMAP = fixed_interval_col_direction_thread_blocking_operator(64, pad) +
 * thread_total_reduce_operator + warp_bit_map_operator / tblock_thread_bit_map_operator (token_test.cc:1253-1315, :1518-1582).  GS_ERR_PADDING_RATE where
 * the reference's padding assert fires; GS_ERR_UNSUPPORTED for families that are not one of the seven drivers'. */
gs_status gs_transform_device(gs_ctx *ctx, int family, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row_dev, const uint32_t *col_dev, const float *val_dev, int N,
                              int sparse_coarsen, int coarsen, int half, int as_id, gs_plan_desc *desc_out);
/* the merge-path chains on the device: [sort_operator +] merge_path_{thread (level 0), warp (1), tblock (2)}_operator(work_size) + the level's total
 * reduction (get_begin_{rows,nzs}_of_level_after_merge_path.cc: one binary search per part instead of the reference's walk from the first row).
 * desc_out is what compile() gives for such a chain: GS_FAMILY_B200_MERGE over the level's first_nz_indices / first_row_indices_without_ending and
 * GLOBAL nz_row_indices; grid / block / vector_width are the launch-shape hints the chain's grid_block_operator would have recorded.
 * coo_on_host != 0: the three COO pointers are host arrays. */
gs_status gs_transform_device_merge_path(gs_ctx *ctx, int level, uint32_t work_size, int sort, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row_dev, const uint32_t *col_dev,
                                         const float *val_dev, int coo_on_host, int N, int coarsen, int vector_width, const uint32_t grid[2], const uint32_t block[2], int half, int as_id,
                                         gs_plan_desc *desc_out);
/* the same from host arrays (H2D, then the device passes); ms_device = the device time of the passes alone (CUDA events), may be NULL */
gs_status gs_transform_host_coo(gs_ctx *ctx, int family, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row, const uint32_t *col, const float *val, int N,
                                int sparse_coarsen, int coarsen, int half, int as_id, gs_plan_desc *desc_out, float *ms_device);
/* a device array of the context back on the host (tests, dumps): len in elements, dtype = its device gs_dtype; host may be NULL to ask for the size */
gs_status gs_download_array(gs_ctx *ctx, const char *key, void *host, uint64_t capacity_bytes, uint64_t *len, int *dtype);

gs_status gs_plan_create(gs_ctx *ctx, const gs_plan_desc *desc, int N, int half, gs_plan **out);
void gs_plan_destroy(gs_plan *p);
/* override the family the plan launches with (e.g. GS_FAMILY_B200_ROW_SMEM on a thread_total chain) */
gs_status gs_plan_set_kernel(gs_plan *p, int family);

/* dense B (K x N row-major, x_arr[col*N+n]; total_BMT...token.cc:872-877): dtype GS_F32 or GS_F16 */
gs_status gs_set_dense(gs_plan *p, const void *B_host, int dtype);
gs_status gs_set_dense_device(gs_plan *p, const void *B_dev);   /* already on the device, plan's value type */
/* enqueue one SpMM (C zeroed first when the family accumulates with atomics) on gs_stream() */
gs_status gs_launch(gs_plan *p);
/* `repeat` launches timed with CUDA events on gs_stream(); ms_per_iter = mean */
gs_status gs_run(gs_plan *p, int repeat, float *ms_per_iter);
/* C (rows x N row-major, y_arr[row*N+n]) in the plan's value type (f32, or f16 when half) */
gs_status gs_get_output(gs_plan *p, void *C_host);
void *gs_output_device(gs_plan *p);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  One SpMM of a plan that runs the
 * GS_FAMILY_B200_TCGEN05 kernel, with the caller's B (K x N fp16, device, 16-byte aligned), the caller's C (rows x N fp16,
 * device, 16-byte aligned) and the caller's stream (cudaStream_t; NULL = the legacy default stream).  The kernel's scratch
 * -- the stream-K workspace and the hand-off flags gs_plan_create allocates per plan -- is a file-scope device array of
 * 5 309 696 bytes that every GPU which loads the library holds until exit; this call allocates and frees nothing (no
 * cudaMalloc, cudaFree or allocating library routine).  Limits, beyond which GS_ERR_INVALID is returned with the limit
 * in gs_last_error(): N <= 64, at most 160 CTAs per launch (one per SM), at most 16 384 row panels (2 097 152 rows).
 * Every call shares the array: calls that overlap on the GPU (different streams, or several host threads) race; calls
 * on one stream are ordered.  Rows of C the matrix does not have (trimmed trailing empty rows) are not written. */
gs_status gs_launch_no_alloc(gs_plan *p, const void *B_dev, void *C_dev, void *stream);
/* launches write C into the caller's device buffer (rows x N, plan's value type, 16-byte aligned) instead of the plan's own; NULL
 * switches back.  Rows the matrix does not have are not written: zero the buffer first.  (How a row block lands in place in a
 * full-size C for the all-gather.) */
gs_status gs_plan_set_output_device(gs_plan *p, void *C_dev);
/* host-buffer round trip the reference's execute_binary does through files (executor.cc:6-58):
 * H2D of B, launch, D2H of C, all on gs_stream(); returns after the stream drained */
gs_status gs_spmm_host(gs_plan *p, const void *B_host, void *C_host);
/* the same round trip without the host wait: H2D on a copy-in stream, launch on one of up to four launch streams
 * (gs_stream() and three more, taken in turn; GS_ASYNC_STREAMS=1..4), D2H on a copy-out stream, chained by events,
 * so that consecutive calls on different plans overlap copies and kernels; calls on one plan stay in order.  Whatever
 * is put on gs_stream() afterwards (gs_launch, gs_get_output ...) is ordered behind them.  B_host and C_host must be
 * pinned and stay valid until gs_synchronize() returns; C_host is complete only then. */
gs_status gs_spmm_host_async(gs_plan *p, const void *B_host, void *C_host);
gs_status gs_synchronize(gs_ctx *ctx); /* drains gs_stream(), the other launch streams and both copy streams */
/* bytes one launch must move by SURVEY.md section 8d: stored nnz * (index + value) + the chain's registered offset / bitmap arrays
 * (+ row permutation) + K*N + M*N values, with the device's dtypes -- the same number whatever kernel the plan launches */
uint64_t gs_plan_algorithmic_bytes(const gs_plan *p);
/* bytes the launched kernel's own layout streams per launch (2:4 image, tile stream, or the chain's arrays): format waste is
 * gs_plan_format_bytes / gs_plan_algorithmic_bytes */
uint64_t gs_plan_format_bytes(const gs_plan *p);
/* bytes of the chain's registered offset arrays that a closed-form model reproduces exactly (what model-driven index compression takes out
 * of the byte stream: first_row_indices of a one-row-per-block level, first_nz_indices of a padded nnz-direction level ...) */
uint64_t gs_plan_model_bytes(const gs_plan *p);
int gs_plan_launch_count(const gs_plan *p); /* kernels per gs_launch (memset excluded) */
const char *gs_plan_kernel_name(const gs_plan *p);
/* GS_FAMILY_B200_SP24 with GS_SP24_PROBE=1 in the environment at plan creation: the 4 metadata columns of the first
 * tile as tensor memory holds them after tcgen05.cp (128 lanes x 4 words), for the layout test */
gs_status gs_plan_sp24_probe(const gs_plan *p, uint32_t *out512);

/* execute_binary(path, matrix, &ms_total, &gflops) of executor.hpp:15 for an uploaded plan:
 * 1 warm-up + `repeat` launches, gflops = FLOAT_RATE*origin_nnz*N*repeat/t/1e9 (code_generator.cc:582-624) */
gs_status gs_execute(gs_plan *p, int repeat, float *ms_total, float *gflops);

/* execute_binary(execute_path, matrix_file, &ms_total, &gflops) of executor.hpp:15 / executor.cc:6-58, on the reference's own hand-off: the
 * directory generate_final_program wrote (the reference's or gsh_session_generate_final_program's) holds one text file per array the
 * kernel reads; the matrix file is read again for the shape and for the host check (code_generator.cc:631-638), N / HALF / VECTOR_WIDTH
 * come from ./global_config.json as in the reference.  No compiler, no fork: the arrays go through gs_upload_array, `kernel` is a
 * gs_family, GS_KERNEL_AUTO or GS_KERNEL_CHAIN.  Prints `correct` or `wrong number:<n>` like the generated program, writes perf_result
 * into the directory, and returns GS_ERR_INTERNAL when the check fails.  `repeat` is used when the directory does not name one. */
gs_status gs_execute_binary(const char *execute_path, const char *matrix_file, int repeat, int kernel, float *ms_total, float *gflops);

/* ------------------------------------------------------------------ device: several GPUs ----- */
/* The reference has no multi-GPU path (executor.hpp:15 hands one program to one device).  North star (3):
 * A is cut into nnz-balanced contiguous row blocks (balanced_interval rule, data_transform_common.cc:934-988), one block per GPU, B is
 * replicated, C rows are disjoint, NCCL only for an optional all-gather of C.
 *
 * One process per GPU (torchrun): every rank owns a gs_ctx; gs_comm_* gives the contexts a communicator.  rank 0 calls
 * gs_comm_unique_id (128 bytes, ncclGetUniqueId), ships the bytes to the other ranks by any means, every rank calls gs_comm_init.
 * gs_comm_allgather_rows: rank g's rows [row_cuts[g], row_cuts[g+1]) already sit at their place in C_full_dev (gs_plan_set_output_device);
 * after the call, enqueued on gs_stream(), every rank holds all rows (ragged blocks: one group of ncclBroadcast).  NCCL is reached
 * through dlopen; GS_ERR_UNSUPPORTED when libnccl.so.2 is absent. */
gs_status gs_comm_unique_id(void *id128);
gs_status gs_comm_init(gs_ctx *ctx, int nranks, int rank, const void *id128);
gs_status gs_comm_allgather_rows(gs_ctx *ctx, void *C_full_dev, const uint64_t *row_cuts, uint64_t row_bytes);
void gs_comm_destroy(gs_ctx *ctx);

/* One process, several GPUs: SURVEY.md section 8b's gs_create(ctx, device_ids, n_dev).  gs_exec_build cuts the matrix of `ms` (as
 * loaded) into gs_exec_devices() row blocks, runs the driver chain `family` (test_spmm_<family>, token_test.cc:1003-1582) on every
 * block and plans it on its device with `kernel` (a gs_family, GS_KERNEL_AUTO or GS_KERNEL_CHAIN).  gs_exec_run: every device runs
 * `repeat` launches on its own stream, enqueued by its own host thread, CUDA-event timed; *ms_per_iter_max = the slowest device.
 * gs_exec_get_output: C (rows x N) in the matrix's row order.  gs_exec_enable_allgather + gs_exec_allgather_output: C replicated on
 * every device (NCCL, in place), *ms = the slowest device's time for the collective; gs_exec_get_replica reads device g's copy. */
typedef struct gs_exec gs_exec;
gs_status gs_exec_create(gs_exec **out, const int *device_ids, int n_dev);
void gs_exec_destroy(gs_exec *ex);
int gs_exec_devices(const gs_exec *ex);
gs_status gs_exec_build(gs_exec *ex, const gsh_metadata *ms, int family, int N, int sparse_coarsen, int coarsen, int half, int kernel);
/* the same partition with every block's chain built by its device (gs_transform_host_coo above, one host thread per device) instead of the host
 * pipeline; families: those gs_transform_device builds */
gs_status gs_exec_build_on_device(gs_exec *ex, const gsh_metadata *ms, int family, int N, int sparse_coarsen, int coarsen, int half, int kernel);
gs_status gs_exec_row_cuts(const gs_exec *ex, uint64_t *cuts /* n_dev + 1 */, uint64_t *block_nnz /* n_dev, may be NULL */);
const char *gs_exec_kernel_name(const gs_exec *ex, int g);
gs_status gs_exec_set_dense(gs_exec *ex, const void *B_host, int dtype);
gs_status gs_exec_run(gs_exec *ex, int repeat, float *ms_per_iter_max, float *ms_per_iter_dev /* n_dev, may be NULL */);
gs_status gs_exec_get_output(gs_exec *ex, void *C_host);
gs_status gs_exec_enable_allgather(gs_exec *ex);
gs_status gs_exec_allgather_output(gs_exec *ex, float *ms);
gs_status gs_exec_get_replica(gs_exec *ex, int g, void *C_host);

#ifdef __cplusplus
}
#endif
#endif /* GSPARSE_B200_H */
