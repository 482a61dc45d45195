// Device-side transform steps: the chains the seven reference drivers build on the host with
// O(nnz) scalar loops over u64 vectors (0.6 s at C2, 5.8 s at C4 against a 12 us / 5 ms kernel) rebuilt as CUDA passes --
// binary searches, block scans, a stable LSD radix sort, warp ballots.  No CUB, no Thrust.  Every array is the one the host
// pipeline (csrc/host/steps.cc, pinned byte for byte to the reference's dumps) produces for the same matrix, in the device dtype
// engine.cu's upload() would give it; tests/test_device_transform_gpu.py compares them byte for byte.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

namespace gsd {

struct built_array {
    std::string pos, name; // "THREAD_META", "first_nz_indices"
    void *ptr = nullptr;   // cudaMalloc'd, len * element bytes + 16
    uint64_t len = 0;
    int dtype = 0;         // gs_dtype
    uint64_t max_value = 0, align = 0;
    bool iota = false;
    bool linear = false;   // entry i == i * linear_step (the only closed form the kernels use: arr_optimization.hpp LINEAR_COMPRESS)
    uint64_t linear_step = 0;
};

struct transform_out {
    std::vector<built_array> arrays;
    uint64_t stored_nnz = 0; // after padding
    uint64_t rows_kept = 0;  // rows the chain's row level spans (trailing empty rows trimmed by the sort)
    bool has_perm = false;
    std::string error;       // set when the chain refuses the matrix (padding rate, unsorted input ...)
    int refused = 0;         // 1: the reference's padding-rate refusal (GS_ERR_PADDING)
    void release();          // frees every array still owned
};

// COO on the device: row-major order (row ids ascending), u32 ids, fp32 values.
struct coo_dev {
    const uint32_t *row = nullptr, *col = nullptr;
    const float *val = nullptr;
    uint64_t rows = 0, cols = 0, nnz = 0;
};

// sort_operator (optional) + fixed_interval_row_direction_thread_blocking_operator(1, pad every row to a multiple of S):
// GLOBAL original_nz_row_indices, nz_{col_indices,vals,row_indices}; `level` ("THREAD_META" / "WARP_META" / "TBLOCK_META") first_row_indices, first_nz_indices
cudaError_t transform_row_chain(const coo_dev &m, bool sort, uint32_t S, bool half, const char *level, double pad_bound, cudaStream_t s, transform_out &out);

// fixed_interval_nnz_direction_thread_blocking_operator(32, pad) + thread_bit_map_operator(THREAD_META, size):
// nz_* padded to a multiple of 32; THREAD first_row_indices, first_nz_indices, thread_bit_map, segment_ptr, segment_empty_flag (packed
// `size` BMTs per word), segment_empty_row_indices
// warp_level: the warp_segment driver chain -- thread_bit_map_operator(WARP_META, size) forces a segment start at every `size`-th BMT, and
// warp_segment_reduce_operator adds THREAD segment_offset and WARP first_BMT_indices (BMWs of `size` BMTs)
cudaError_t transform_nnz_chain(const coo_dev &m, uint32_t size, bool warp_level, bool half, double pad_bound, cudaStream_t s, transform_out &out);

// fixed_interval_col_direction_thread_blocking_operator(col_block, pad) + thread_total_reduce_operator + warp_bit_map_operator (warp_level: BMWs of
// `group` BMTs) or tblock_thread_bit_map_operator (BMTBs of `group` BMTs): the warp_bitmap / block_bitmap driver chains
cudaError_t transform_col_chain(const coo_dev &m, uint32_t col_block, uint32_t group, bool warp_level, bool half, double pad_bound, cudaStream_t s, transform_out &out);

// [sort_operator +] merge_path_{thread,warp,tblock}_operator(work): the (sorted) stream, `level` first_row_indices_without_ending and first_nz_indices
cudaError_t transform_merge_path_chain(const coo_dev &m, bool sort, uint32_t work, const char *level, bool half, double pad_bound, cudaStream_t s, transform_out &out);

// pieces the other plan-time builders reuse: exclusive scan (out has n + 1 entries, in == out allowed) and CSR offsets of ascending row ids
cudaError_t scan_exclusive_u32(const uint32_t *in, uint32_t *out, uint64_t n, cudaStream_t s);
cudaError_t rowptr_from_sorted_rows(const uint32_t *row, uint32_t nnz, uint32_t rows, uint32_t *rowptr, cudaStream_t s);

} // namespace gsd
