// The seven kernel families of the reference driver as hand-written, templated CUDA kernels.
// Each kernel consumes exactly the metadata arrays the reference's reduction token of the same
// family registers as kernel inputs (reduction_token/*.cc; pinned by tests/golden "registered"
// lists) and keeps the reference's thread mapping: one dense column group per thread along one
// block dimension, BMT / lane along the other (token_test.cc launch shapes).
// B is K x N row-major (x_arr[col*N+n]), C is M x N row-major (y_arr[row*N+n]); products are
// accumulated in fp32 whatever the storage type (HALF: fp16 storage, fp32 accumulate).
//
// These kernels are the parity-complete coverage of the operator chains; the throughput path
// for the pruned-weight configs is in kernels_b200.cu.
#include "common.cuh"

#include "../../../include/gsparse_b200.h"

namespace gsd {

namespace {

constexpr int MAXC = 8; // largest coarsen factor (dense columns per thread)

template <typename ColT, typename ValT> struct mat_view {
    const ColT *col;
    const ValT *val;
    const ValT *B;
    ValT *C;
    int N;
};

template <typename ColT, typename ValT> __device__ __forceinline__ mat_view<ColT, ValT> view(const launch_params &p)
{
    return {(const ColT *)p.a.col, (const ValT *)p.a.val, (const ValT *)p.B, (ValT *)p.C, p.N};
}

__device__ __forceinline__ uint32_t out_row(const dev_arrays &a, uint32_t row) { return a.perm ? a.perm[row] + a.begin_row : row + a.begin_row; }

// one non-zero times `coarsen` dense columns starting at n0
template <typename ColT, typename ValT>
__device__ __forceinline__ void fma_nz(const mat_view<ColT, ValT> &m, uint64_t nz, int n0, int coarsen, float (&acc)[MAXC])
{
    float v = to_f32(m.val[nz]);
    const ValT *b = m.B + (size_t)m.col[nz] * m.N + n0;
#pragma unroll
    for (int f = 0; f < MAXC; f++)
        if (f < coarsen && n0 + f < m.N) acc[f] += v * to_f32(b[f]);
}

template <typename ValT> __device__ __forceinline__ void store_row(ValT *C, int N, uint32_t row, int n0, int coarsen, const float (&acc)[MAXC])
{
#pragma unroll
    for (int f = 0; f < MAXC; f++)
        if (f < coarsen && n0 + f < N) C[(size_t)row * N + n0 + f] = from_f32<ValT>(acc[f]);
}

template <typename ValT> __device__ __forceinline__ void add_row(ValT *C, int N, uint32_t row, int n0, int coarsen, const float (&acc)[MAXC])
{
#pragma unroll
    for (int f = 0; f < MAXC; f++)
        if (f < coarsen && n0 + f < N && acc[f] != 0.0f) atomic_add_out(C, (size_t)row * N + n0 + f, acc[f]);
}

// ---- thread-total: one BMT (row block) per thread, direct store (rows disjoint) ------------------
// total_BMT_result_reduce_to_one_register_token.cc:470-913; loop vars code_generator.cc:1563-1571
template <typename ColT, typename ValT> __global__ void k_thread_total(launch_params p)
{
    auto m = view<ColT, ValT>(p);
    int n0 = (blockIdx.y * blockDim.x + threadIdx.x) * p.coarsen;
    if (n0 >= p.N) return;
    for (uint32_t b = blockIdx.x * blockDim.y + threadIdx.y; b < p.a.n_t; b += gridDim.x * blockDim.y) {
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t lo = p.a.t_first_nz[b], hi = p.a.t_first_nz[b + 1];
        for (uint32_t nz = lo; nz < hi; nz++) fma_nz(m, nz, n0, p.coarsen, acc);
        // a BMT of several rows cannot be reduced to one register per column: the chain guarantees row block 1
        store_row(m.C, p.N, out_row(p.a, p.a.t_first_row[b]), n0, p.coarsen, acc);
    }
}

// ---- warp-total: one BMW (row) per warp, lanes stride the non-zeros, shuffle tree ----------------
// total_warp_result_reduce_to_one_register_token.cc:525-840.  block = (32 lanes, dense columns)
template <typename ColT, typename ValT> __global__ void k_warp_total(launch_params p)
{
    auto m = view<ColT, ValT>(p);
    int n0 = (blockIdx.y * blockDim.y + threadIdx.y) * p.coarsen;
    bool live = n0 < p.N;
    for (uint32_t w = blockIdx.x; w < p.a.n_w; w += gridDim.x) {
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t lo = p.a.w_first_nz[w], hi = p.a.w_first_nz[w + 1];
        if (live)
            for (uint32_t nz = lo + threadIdx.x; nz < hi; nz += 32) fma_nz(m, nz, n0, p.coarsen, acc);
#pragma unroll
        for (int f = 0; f < MAXC; f++)
            if (f < p.coarsen) {
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) acc[f] += __shfl_down_sync(0xffffffffu, acc[f], d);
            }
        if (threadIdx.x == 0 && live) store_row(m.C, p.N, out_row(p.a, p.a.w_first_row[w]), n0, p.coarsen, acc);
    }
}

// ---- tblock-total: one BMTB (row) per CTA, shared-memory tree over threadIdx.y -------------------
// total_block_reduce_to_one_register_token.cc:374-525.  block = (dense columns, nnz lanes)
template <typename ColT, typename ValT> __global__ void k_tblock_total(launch_params p)
{
    extern __shared__ float red[]; // blockDim.y * blockDim.x * coarsen
    auto m = view<ColT, ValT>(p);
    int n0 = (blockIdx.y * blockDim.x + threadIdx.x) * p.coarsen;
    bool live = n0 < p.N;
    for (uint32_t t = blockIdx.x; t < p.a.n_tb; t += gridDim.x) {
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t lo = p.a.tb_first_nz[t], hi = p.a.tb_first_nz[t + 1];
        if (live)
            for (uint32_t nz = lo + threadIdx.y; nz < hi; nz += blockDim.y) fma_nz(m, nz, n0, p.coarsen, acc);
        for (int f = 0; f < p.coarsen; f++) red[(threadIdx.y * blockDim.x + threadIdx.x) * p.coarsen + f] = acc[f];
        __syncthreads();
        for (int s = 1; s < (int)blockDim.y; s <<= 1) { // log-step tree; blockDim.y need not be a power of two
            if ((threadIdx.y % (2 * s)) == 0 && threadIdx.y + s < blockDim.y)
                for (int f = 0; f < p.coarsen; f++)
                    red[(threadIdx.y * blockDim.x + threadIdx.x) * p.coarsen + f] += red[((threadIdx.y + s) * blockDim.x + threadIdx.x) * p.coarsen + f];
            __syncthreads();
        }
        if (threadIdx.y == 0 && live) {
            for (int f = 0; f < p.coarsen; f++) acc[f] = red[threadIdx.x * p.coarsen + f];
            store_row(m.C, p.N, out_row(p.a, p.a.tb_first_row[t]), n0, p.coarsen, acc);
        }
        __syncthreads();
    }
}

// segmented sum over the 32 lanes of a warp: `heads` bit i = lane i starts a segment; the sum of a
// segment lands on its head lane (replaces the reference's 5-step compare-row shuffle ladder)
__device__ __forceinline__ float warp_segmented_sum(float v, uint32_t heads, int lane)
{
    uint32_t above = heads & ~((2u << lane) - 1u); // heads strictly after this lane
    int seg_end = above ? __ffs(above) - 1 : 32;   // first lane of the next segment
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float o = __shfl_down_sync(0xffffffffu, v, d);
        if (lane + d < seg_end) v += o;
    }
    return v;
}

// ---- warp-bitmap: BMT = row chunk (<= 64 nnz); per-BMT partial, segmented warp sum by the
// bit_map_of_thread head bits, heads atomicAdd.  warp_bit_map_reduce_token.cc:94-392.
// block = (32 lanes = 32 consecutive BMTs, dense columns)
template <typename ColT, typename ValT> __global__ void k_warp_bitmap(launch_params p)
{
    auto m = view<ColT, ValT>(p);
    int lane = threadIdx.x;
    int n0 = (blockIdx.y * blockDim.y + threadIdx.y) * p.coarsen;
    bool live = n0 < p.N;
    const int vw = p.vector_width;
    uint32_t n_groups = (p.a.n_t + 31) / 32;
    for (uint32_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
        uint32_t b = g * 32 + lane;
        bool has = b < p.a.n_t;
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t row = 0;
        bool head = false;
        if (has) {
            uint32_t w = b / vw, k = b % vw; // BMW of this BMT and its position inside it
            // first_nz_indices_relative_to_BMW + the BMW's begin (the chain registers the relative array)
            uint32_t lo, hi;
            if (p.a.t_nz_relative) {
                uint32_t base = p.a.w_first_nz[w];
                lo = base + p.a.t_first_nz[b];
                hi = (k + 1 < (uint32_t)vw && b + 1 < p.a.n_t) ? base + p.a.t_first_nz[b + 1] : p.a.w_first_nz[w + 1];
            } else {
                lo = p.a.t_first_nz[b], hi = p.a.t_first_nz[b + 1];
            }
            if (live)
                for (uint32_t nz = lo; nz < hi; nz++) fma_nz(m, nz, n0, p.coarsen, acc);
            row = p.a.t_first_row_we[b];
            head = (p.a.w_bit_map[w] >> k) & 1ull;
        }
        uint32_t heads = __ballot_sync(0xffffffffu, head || lane == 0 || !has);
#pragma unroll
        for (int f = 0; f < MAXC; f++)
            if (f < p.coarsen) acc[f] = warp_segmented_sum(acc[f], heads, lane);
        if (has && live && ((heads >> lane) & 1u)) add_row(m.C, p.N, out_row(p.a, row), n0, p.coarsen, acc);
    }
}

// ---- tblock-bitmap: per-BMT partials to shared memory, heads (bit_map_of_thread) add the
// segment_offset following partials and atomicAdd.  tblock_bit_map_reduce_token.cc:83-189.
// block = (dense columns, block_size BMTs).  The reference's segment_offset leaves the run that
// reaches the end of the array at 0 (segment_offset.cc), which would drop the last row's trailing
// chunks; a head with offset 0 therefore extends over directly following non-head BMTs of its CTA.
template <typename ColT, typename ValT> __global__ void k_tblock_bitmap(launch_params p)
{
    extern __shared__ float part[]; // blockDim.y * blockDim.x * coarsen
    auto m = view<ColT, ValT>(p);
    int n0 = (blockIdx.y * blockDim.x + threadIdx.x) * p.coarsen;
    bool live = n0 < p.N;
    for (uint32_t t = blockIdx.x; t < p.a.n_tb; t += gridDim.x) {
        uint32_t b0 = p.a.tb_first_bmt[t], b1 = p.a.tb_first_bmt[t + 1];
        uint32_t b = b0 + threadIdx.y;
        bool has = b < b1;
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (has && live) {
            uint32_t lo = p.a.t_first_nz[b], hi = p.a.t_first_nz[b + 1];
            for (uint32_t nz = lo; nz < hi; nz++) fma_nz(m, nz, n0, p.coarsen, acc);
        }
        for (int f = 0; f < p.coarsen; f++) part[(threadIdx.y * blockDim.x + threadIdx.x) * p.coarsen + f] = acc[f];
        __syncthreads();
        if (has && live && p.a.t_head[b]) {
            uint32_t follow = p.a.t_seg_offset[b];
            if (follow == 0)
                while (b + follow + 1 < b1 && p.a.t_head[b + follow + 1] == 0) follow++;
            for (uint32_t j = 1; j <= follow && b + j < b1; j++)
                for (int f = 0; f < p.coarsen; f++) acc[f] += part[((threadIdx.y + j) * blockDim.x + threadIdx.x) * p.coarsen + f];
            add_row(m.C, p.N, out_row(p.a, p.a.t_first_row_we[b]), n0, p.coarsen, acc);
        }
        __syncthreads();
    }
}

// row of segment `seg` of BMT b: consecutive rows unless the BMT spans empty rows (segment_empty_flag),
// in which case segment_empty_row_indices holds the offset from the BMT's first row
__device__ __forceinline__ uint32_t seg_row(const dev_arrays &a, uint32_t b, uint32_t first_row, uint32_t seg, bool gaps)
{
    return gaps ? first_row + a.t_seg_empty_rows[a.t_seg_ptr[b] + seg] : first_row + seg;
}

// ---- thread-bitmap: fixed nnz per BMT regardless of rows; walk the row-start bit map, flush at each
// row boundary with atomicAdd.  thread_bit_map_reduce_to_two_register_token.cc:819-1616.
// block = (dense columns, BMTs)
template <typename ColT, typename ValT> __global__ void k_thread_bitmap(launch_params p)
{
    auto m = view<ColT, ValT>(p);
    int n0 = (blockIdx.y * blockDim.x + threadIdx.x) * p.coarsen;
    if (n0 >= p.N) return;
    for (uint32_t b = blockIdx.x * blockDim.y + threadIdx.y; b < p.a.n_t; b += gridDim.x * blockDim.y) {
        uint32_t lo = p.a.t_first_nz[b], hi = p.a.t_first_nz[b + 1];
        uint64_t map = p.a.t_bit_map[b];
        bool gaps = (p.a.t_seg_empty_flag[b / p.vector_width] >> (b % p.vector_width)) & 1ull;
        uint32_t first_row = p.a.t_first_row[b], row = first_row, seg = 0;
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (uint32_t k = 0; lo + k < hi; k++) {
            if (k > 0 && ((map >> k) & 1ull)) {
                add_row(m.C, p.N, out_row(p.a, row), n0, p.coarsen, acc);
#pragma unroll
                for (int f = 0; f < MAXC; f++) acc[f] = 0.0f;
                seg++;
                row = seg_row(p.a, b, first_row, seg, gaps);
            }
            fma_nz(m, lo + k, n0, p.coarsen, acc);
        }
        add_row(m.C, p.N, out_row(p.a, row), n0, p.coarsen, acc);
    }
}

// ---- warp-segment: thread-bitmap walk, but the partial before the first row start ("cont") and the
// partial after the last one ("tail") are combined across the lanes of a warp before touching C.
// warp_segment_reduce_token.cc:94-440 does this with a shfl_up scan + shfl_down by segment_offset; here
// one segmented suffix sum over ballot masks.  block = (32 lanes = 32 consecutive BMTs, dense columns)
template <typename ColT, typename ValT> __global__ void k_warp_segment(launch_params p)
{
    auto m = view<ColT, ValT>(p);
    int lane = threadIdx.x;
    int n0 = (blockIdx.y * blockDim.y + threadIdx.y) * p.coarsen;
    bool live = n0 < p.N;
    uint32_t n_groups = (p.a.n_t + 31) / 32;
    for (uint32_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
        uint32_t b = g * 32 + lane;
        bool has = b < p.a.n_t;
        float acc[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0}, cont[MAXC] = {0, 0, 0, 0, 0, 0, 0, 0};
        bool started = false; // this BMT contains a row start
        uint32_t first_row = 0, row = 0;
        if (has) {
            uint32_t lo = p.a.t_first_nz[b], hi = p.a.t_first_nz[b + 1];
            uint64_t map = p.a.t_bit_map[b];
            bool gaps = (p.a.t_seg_empty_flag[b / p.vector_width] >> (b % p.vector_width)) & 1ull;
            first_row = p.a.t_first_row[b];
            row = first_row;
            uint32_t seg = 0;
            for (uint32_t k = 0; lo + k < hi; k++) {
                if ((map >> k) & 1ull) {
                    if (!started) {
                        started = true;
#pragma unroll
                        for (int f = 0; f < MAXC; f++) cont[f] = acc[f]; // belongs to the row an earlier BMT started
                    } else if (live) {
                        add_row(m.C, p.N, out_row(p.a, row), n0, p.coarsen, acc); // a row wholly inside this BMT
                    }
                    if (k > 0) {
                        seg++;
                        row = seg_row(p.a, b, first_row, seg, gaps);
                    }
#pragma unroll
                    for (int f = 0; f < MAXC; f++) acc[f] = 0.0f;
                }
                if (live) fma_nz(m, lo + k, n0, p.coarsen, acc);
            }
            if (!started) {
#pragma unroll
                for (int f = 0; f < MAXC; f++) cont[f] = acc[f], acc[f] = 0.0f;
            }
        }
        // runs of lanes end at a lane that contains a row start; suffix-sum cont inside the runs
        uint32_t flagged = __ballot_sync(0xffffffffu, started);
        uint32_t at_or_after = flagged >> lane;
        int run_end = at_or_after ? lane + __ffs(at_or_after) - 1 : 31;
        bool before_first = (flagged & ((1u << lane) - 1u)) == 0; // no row start in an earlier lane: cont goes straight to C
#pragma unroll
        for (int f = 0; f < MAXC; f++) {
            if (f >= p.coarsen) continue;
            float x = cont[f];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                float o = __shfl_down_sync(0xffffffffu, x, d);
                if (lane + d <= run_end) x += o;
            }
            float next = __shfl_down_sync(0xffffffffu, x, 1);
            if (lane == 31) next = 0.0f;
            acc[f] += started ? next : 0.0f; // tail + everything later lanes contribute to the same row
        }
        if (has && live) {
            if (started) add_row(m.C, p.N, out_row(p.a, row), n0, p.coarsen, acc);
            if (before_first) add_row(m.C, p.N, out_row(p.a, first_row), n0, p.coarsen, cont);
        }
    }
}

template <typename ColT, typename ValT> cudaError_t launch_typed(int family, const launch_params &p, cudaStream_t s)
{
    const int c = p.coarsen;
    const int ncols = (p.N + c - 1) / c; // dense column groups
    int dx = ncols < 32 ? ncols : 32;    // threads along the dense dimension
    unsigned gy = (ncols + dx - 1) / dx;
    switch (family) {
    case GS_FAMILY_THREAD_TOTAL: {
        int dy = 128 / dx > 0 ? 128 / dx : 1;
        unsigned gx = (p.a.n_t + dy - 1) / dy;
        k_thread_total<ColT, ValT><<<dim3(gx ? gx : 1, gy), dim3(dx, dy), 0, s>>>(p);
        break;
    }
    case GS_FAMILY_THREAD_BITMAP: {
        int dy = 128 / dx > 0 ? 128 / dx : 1;
        unsigned gx = (p.a.n_t + dy - 1) / dy;
        k_thread_bitmap<ColT, ValT><<<dim3(gx ? gx : 1, gy), dim3(dx, dy), 0, s>>>(p);
        break;
    }
    case GS_FAMILY_WARP_TOTAL:
        k_warp_total<ColT, ValT><<<dim3(p.a.n_w < 65535u * 32u ? p.a.n_w : 65535u * 32u, gy), dim3(32, dx), 0, s>>>(p);
        break;
    case GS_FAMILY_WARP_BITMAP: {
        unsigned groups = (p.a.n_t + 31) / 32;
        k_warp_bitmap<ColT, ValT><<<dim3(groups ? groups : 1, gy), dim3(32, dx), 0, s>>>(p);
        break;
    }
    case GS_FAMILY_WARP_SEGMENT: {
        unsigned groups = (p.a.n_t + 31) / 32;
        k_warp_segment<ColT, ValT><<<dim3(groups ? groups : 1, gy), dim3(32, dx), 0, s>>>(p);
        break;
    }
    case GS_FAMILY_TBLOCK_TOTAL: {
        int dy = 256 / dx;
        size_t smem = (size_t)dx * dy * c * sizeof(float);
        k_tblock_total<ColT, ValT><<<dim3(p.a.n_tb, gy), dim3(dx, dy), smem, s>>>(p);
        break;
    }
    case GS_FAMILY_TBLOCK_BITMAP: {
        int dy = p.block_size;
        if (dx * dy > 1024) return cudaErrorInvalidConfiguration;
        size_t smem = (size_t)dx * dy * c * sizeof(float);
        k_tblock_bitmap<ColT, ValT><<<dim3(p.a.n_tb, gy), dim3(dx, dy), smem, s>>>(p);
        break;
    }
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_reference_family(int family, const launch_params &p, cudaStream_t s)
{
    if (p.coarsen < 1 || p.coarsen > MAXC) return cudaErrorInvalidValue;
    if (p.a.col_is_u16) return p.a.val_is_f16 ? launch_typed<uint16_t, __half>(family, p, s) : launch_typed<uint16_t, float>(family, p, s);
    return p.a.val_is_f16 ? launch_typed<uint32_t, __half>(family, p, s) : launch_typed<uint32_t, float>(family, p, s);
}

const char *reference_family_kernel_name(int family)
{
    static const char *n[] = {"k_thread_total", "k_warp_total", "k_warp_bitmap", "k_thread_bitmap", "k_warp_segment", "k_tblock_total", "k_tblock_bitmap"};
    return (family >= 0 && family < 7) ? n[family] : "unknown";
}

} // namespace gsd
