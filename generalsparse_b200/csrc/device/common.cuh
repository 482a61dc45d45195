// Shared device-side definitions of the SpMM engine (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsd {

// canonical device arrays of one sub-matrix (widened at upload, see engine.cu)
struct dev_arrays {
    const void *col = nullptr;  // u16 or u32
    const void *val = nullptr;  // f32 or f16
    int col_is_u16 = 0;
    int val_is_f16 = 0;
    const uint32_t *perm = nullptr;          // GLOBAL original_nz_row_indices (sorted row -> original row) or null
    const uint32_t *tb_first_row = nullptr;  // TBLOCK first_row_indices
    const uint32_t *tb_first_nz = nullptr;   // TBLOCK first_nz_indices
    const uint32_t *tb_first_bmt = nullptr;  // TBLOCK first_BMT_indices
    const uint32_t *w_first_row = nullptr;   // WARP first_row_indices
    const uint32_t *w_first_nz = nullptr;    // WARP first_nz_indices
    const uint32_t *w_first_bmt = nullptr;   // WARP first_BMT_indices
    const uint64_t *w_bit_map = nullptr;     // WARP bit_map_of_thread (one word per VECTOR_WIDTH BMTs)
    const uint32_t *t_first_row = nullptr;   // THREAD first_row_indices (with sentinel)
    const uint32_t *t_first_row_we = nullptr;// THREAD first_row_indices_without_ending
    const uint32_t *t_first_nz = nullptr;    // THREAD first_nz_indices (or ..._relative_to_BMW when t_nz_relative)
    int t_nz_relative = 0;
    const uint64_t *t_bit_map = nullptr;     // THREAD thread_bit_map
    const uint32_t *t_head = nullptr;        // THREAD bit_map_of_thread (0/1 per BMT, tblock mode)
    const uint32_t *t_seg_ptr = nullptr;     // THREAD segment_ptr
    const uint32_t *t_seg_offset = nullptr;  // THREAD segment_offset
    const uint64_t *t_seg_empty_flag = nullptr;     // THREAD segment_empty_flag (packed per `size` BMTs)
    const uint32_t *t_seg_empty_rows = nullptr;     // THREAD segment_empty_row_indices
    uint32_t n_tb = 0, n_w = 0, n_t = 0;     // block counts at each level (entries - 1 where a sentinel exists)
    uint64_t nnz = 0;                        // stored (padded) non-zeros
    uint32_t rows = 0, cols = 0;             // C rows / K
    uint32_t begin_row = 0;                  // begin_row_index of the sub-matrix
};

struct launch_params {
    dev_arrays a;
    const void *B = nullptr; // K x N row-major, value type
    void *C = nullptr;       // rows x N row-major, value type
    int N = 0;
    int coarsen = 1;
    int vector_width = 32;   // VECTOR_WIDTH of the chain (BMTs per BMW / flag packing size)
    int block_size = 32;     // BMTs per BMTB in tblock-bitmap mode
    uint32_t first_nz_step = 0; // != 0: THREAD first_nz_indices[i] == first_nz_step * i (model-driven index compression): computed, not loaded
    bool pdl = false;        // programmatic dependent launch allowed (engine.cu)
    bool dep_wait = false;   // B (or whatever else the launch reads) may belong to the operation before it on the stream: a kernel that cannot wait
                             // for it on the device (griddepcontrol.wait before its first read of B) must be launched without the attribute
};

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// accumulate into C: fp32 atomics for float, CAS-free half2-unaligned-safe path for half
__device__ __forceinline__ void atomic_add_out(float *C, size_t idx, float v) { atomicAdd(C + idx, v); }
__device__ __forceinline__ void atomic_add_out(__half *C, size_t idx, float v) { atomicAdd(C + idx, __float2half_rn(v)); }

// cache-hinted streaming loads for data read exactly once (col/val streams)
__device__ __forceinline__ uint4 ld_stream_v4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

} // namespace gsd

// launchers (host functions, defined in kernels_ref.cu / kernels_b200.cu); return cudaError_t of the launch
namespace gsd {
cudaError_t launch_reference_family(int family, const launch_params &p, cudaStream_t s);
const char *reference_family_kernel_name(int family);

struct b200_row_plan; // opaque, kernels_b200.cu
cudaError_t b200_row_prepare(const launch_params &p, int sm_count, b200_row_plan **out, cudaStream_t s);
cudaError_t b200_row_launch(const b200_row_plan *plan, const launch_params &p, cudaStream_t s);
void b200_row_free(b200_row_plan *plan);
const char *b200_row_kernel_name(const b200_row_plan *plan);

struct b200_tile_plan; // opaque, kernels_tile.cu
cudaError_t b200_tile_prepare(const launch_params &p, int sm_count, b200_tile_plan **out, cudaStream_t s);
cudaError_t b200_tile_launch(b200_tile_plan *plan, const launch_params &p, cudaStream_t s);
void b200_tile_free(b200_tile_plan *plan);
const char *b200_tile_kernel_name(const b200_tile_plan *plan);
uint64_t b200_tile_bytes(const b200_tile_plan *plan, uint64_t rows, uint64_t cols);
uint64_t b200_tile_pairs(const b200_tile_plan *plan);
// the launch of gs_launch_no_alloc: workspace and flags in a file-scope device array instead of the plan's allocations
cudaError_t b200_tile_launch_no_alloc(const b200_tile_plan *plan, int device, const void *B, void *C, cudaStream_t s, const char **why);
size_t b200_tile_no_alloc_scratch_bytes();

struct b200_sp24_plan; // opaque, kernels_sp24.cu
cudaError_t b200_sp24_prepare(const launch_params &p, int sm_count, bool allow_remainder, b200_sp24_plan **out, cudaStream_t s);
cudaError_t b200_sp24_launch(b200_sp24_plan *plan, const launch_params &p, cudaStream_t s);
void b200_sp24_free(b200_sp24_plan *plan);
const char *b200_sp24_kernel_name(const b200_sp24_plan *plan);
uint64_t b200_sp24_bytes(const b200_sp24_plan *plan, uint64_t rows, uint64_t cols);
uint64_t b200_sp24_remainder(const b200_sp24_plan *plan);
int b200_sp24_launches(const b200_sp24_plan *plan);
cudaError_t b200_sp24_probe(const b200_sp24_plan *plan, uint32_t *out512);

// kernels_nnz.cu: warp-per-run kernel over the nnz-direction bit-map chains (no plan-time state)
bool b200_nnz_supported(const launch_params &p);
struct b200_nnz_plan; // hot-column copy of the ids for matrices whose B does not fit L2 (may stay null)
cudaError_t b200_nnz_prepare(const launch_params &p, b200_nnz_plan **out, cudaStream_t s);
void b200_nnz_free(b200_nnz_plan *plan);
cudaError_t b200_nnz_launch(const launch_params &p, const b200_nnz_plan *plan, cudaStream_t s);
const char *b200_nnz_kernel_name(const launch_params &p, const b200_nnz_plan *plan);

// kernels_merge.cu: range kernel over merge-path / balanced / column-direction chains
struct b200_merge_plan;
bool b200_merge_supported(int N);
cudaError_t b200_merge_prepare(const launch_params &p, const uint32_t *parts_dev, uint32_t n_parts, const uint32_t *nz_rows_dev, const uint32_t *part_rows_dev, bool parts_are_rows,
                               b200_merge_plan **out, cudaStream_t s);
cudaError_t b200_merge_launch(const b200_merge_plan *plan, const launch_params &p, cudaStream_t s);
void b200_merge_free(b200_merge_plan *plan);
const char *b200_merge_kernel_name(const b200_merge_plan *plan);
bool b200_merge_needs_zero(const b200_merge_plan *plan); // false: the kernel writes every row itself (ticket mode), no memset of C
uint64_t b200_merge_bytes(const b200_merge_plan *plan, uint64_t nnz, uint64_t idx_bytes, uint64_t val_bytes, uint64_t rows, uint64_t cols);
} // namespace gsd
