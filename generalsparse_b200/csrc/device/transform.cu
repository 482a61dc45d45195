// Device-side transform steps (transform.hpp).  The reference files each pass restates are named where the pass is.
#include "transform.hpp"

#include "../../../include/gsparse_b200.h"

#include <cuda_fp16.h>

#include <algorithm>
#include <cstdio>

namespace gsd {

void transform_out::release()
{
    for (built_array &a : arrays) cudaFree(a.ptr), a.ptr = nullptr;
    arrays.clear();
}

namespace {

#define TRY(x)                                                                                                                                       \
    do {                                                                                                                                             \
        cudaError_t e_ = (x);                                                                                                                        \
        if (e_ != cudaSuccess) return e_;                                                                                                            \
    } while (0)

// Every buffer of a transform comes from the device's stream-ordered pool (cudaMallocAsync on the transform's stream) with the pool told to keep
// what is freed: a chain takes ~16 allocations, and plain cudaMalloc / cudaFree cost tens of milliseconds each once peer access is on (a 2-GPU
// bench line measured 465 ms for the passes that take 12 ms on one GPU).  Arrays handed to the context are freed by it with cudaFree, which takes
// pool memory too.
cudaError_t pool_alloc(void **p, uint64_t bytes, cudaStream_t s)
{
    static bool keep[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64 && !keep[dev]) {
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t all = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &all);
        }
        keep[dev] = true;
    }
    return cudaMallocAsync(p, bytes, s);
}

// scratch that lives for one transform
struct scratch {
    std::vector<void *> bufs;
    cudaStream_t stream = nullptr;
    explicit scratch(cudaStream_t s) : stream(s) {}
    ~scratch()
    {
        for (void *p : bufs) cudaFreeAsync(p, stream);
    }
    template <typename T> cudaError_t get(T **p, uint64_t n)
    {
        void *q = nullptr;
        cudaError_t e = pool_alloc(&q, std::max<uint64_t>(n, 1) * sizeof(T) + 16, stream);
        if (e == cudaSuccess) bufs.push_back(q), *p = (T *)q;
        return e;
    }
};

inline unsigned blocks_for(uint64_t n, unsigned per) { return (unsigned)std::max<uint64_t>(1, (n + per - 1) / per); }

// ---- exclusive scan (u32): 4096 entries per CTA, CTA sums scanned by a recursive call ------------------------------------------------------
constexpr int SCAN_T = 512, SCAN_I = 8, SCAN_TILE = SCAN_T * SCAN_I;

__global__ void __launch_bounds__(SCAN_T) k_scan_tile(const uint32_t *in, uint32_t *out, uint32_t *tile_sum, uint64_t n, uint64_t n_out)
{
    __shared__ uint32_t warp_sum[SCAN_T / 32];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_I;
    uint32_t v[SCAN_I], mine = 0;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) v[i] = base + i < n ? in[base + i] : 0u, mine += v[i];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sum[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t s = lane < SCAN_T / 32 ? warp_sum[lane] : 0u, si = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, si, d);
            if (lane >= d) si += t;
        }
        if (lane < SCAN_T / 32) warp_sum[lane] = si - s; // exclusive over warps
        if (lane == SCAN_T / 32 - 1 && tile_sum) tile_sum[blockIdx.x] = si;
    }
    __syncthreads();
    uint32_t run = warp_sum[w] + inc - mine;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) {
        if (base + i < n_out) out[base + i] = run;
        run += v[i];
    }
}

__global__ void __launch_bounds__(SCAN_T) k_scan_add(uint32_t *out, const uint32_t *tile_off, uint64_t n)
{
    const uint32_t off = tile_off[blockIdx.x];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    for (int i = threadIdx.x; i < SCAN_TILE; i += SCAN_T)
        if (base + i < n) out[base + i] += off;
}

// out[i] = in[0] + ... + in[i - 1] for i in [0, n]; out has n + 1 entries (out[n] = the total).  in == out is allowed.
cudaError_t scan_exclusive(const uint32_t *in, uint32_t *out, uint64_t n, scratch &sc, cudaStream_t s)
{
    const uint64_t m = n + 1; // the extra entry reads as zero and receives the total
    const unsigned tiles = blocks_for(m, SCAN_TILE);
    uint32_t *sums = nullptr;
    if (tiles > 1) TRY(sc.get(&sums, (uint64_t)tiles + 1));
    k_scan_tile<<<tiles, SCAN_T, 0, s>>>(in, out, sums, n, m); // reads in[i] for i < n only; entry n receives the total
    TRY(cudaGetLastError());
    if (tiles > 1) {
        TRY(scan_exclusive(sums, sums, tiles, sc, s));
        k_scan_add<<<tiles, SCAN_T, 0, s>>>(out, sums, m);
        TRY(cudaGetLastError());
    }
    return cudaSuccess;
}

// ---- reductions --------------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t gcd32(uint32_t a, uint32_t b)
{
    while (b) {
        const uint32_t t = a % b;
        a = b, b = t;
    }
    return a;
}
struct op_max {
    __device__ static uint32_t f(uint32_t a, uint32_t b) { return a > b ? a : b; }
};
struct op_gcd {
    __device__ static uint32_t f(uint32_t a, uint32_t b) { return gcd32(a, b); }
};
struct op_sum {
    __device__ static uint32_t f(uint32_t a, uint32_t b) { return a + b; }
};
template <typename Op> __global__ void __launch_bounds__(256) k_reduce(const uint32_t *in, uint64_t n, uint32_t *out)
{
    __shared__ uint32_t sh[256];
    uint32_t acc = 0; // neutral for max, gcd and sum
    for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (uint64_t)gridDim.x * 256) acc = Op::f(acc, in[i]);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = Op::f(sh[threadIdx.x], sh[threadIdx.x + d]);
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = sh[0];
}
template <typename Op> cudaError_t reduce_u32(const uint32_t *in, uint64_t n, uint32_t *result_host, scratch &sc, cudaStream_t s)
{
    const unsigned blocks = std::min(blocks_for(n, 256 * 16), 1024u);
    uint32_t *part = nullptr;
    TRY(sc.get(&part, (uint64_t)blocks + 1));
    k_reduce<Op><<<blocks, 256, 0, s>>>(in, n, part);
    k_reduce<Op><<<1, 256, 0, s>>>(part, blocks, part + blocks);
    TRY(cudaGetLastError());
    TRY(cudaMemcpyAsync(result_host, part + blocks, 4, cudaMemcpyDeviceToHost, s));
    return cudaStreamSynchronize(s);
}

// ---- stable LSD radix sort of (key, value) pairs, 8 bits per pass ----------------------------------------------------------------------------
// get_row_order_by_length.cc:67-90 is a counting sort over the row lengths, longest first, stable inside a length class; on the device
// the same order comes out of a stable sort on key = max_len - len.  A CTA owns 2048 consecutive pairs; a warp owns 256 consecutive ones
// and walks them 32 at a time, so "earlier in the input" is (CTA, warp, round, lane) order and the ranks below keep it.
constexpr int RS_WARPS = 8, RS_ROUNDS = 8, RS_TILE = RS_WARPS * 32 * RS_ROUNDS;

__global__ void __launch_bounds__(RS_WARPS * 32) k_radix_hist(const uint32_t *key, uint64_t n, int shift, uint32_t *hist /* [256][tiles] */, uint32_t tiles)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * RS_TILE;
    for (int i = threadIdx.x; i < RS_TILE; i += RS_WARPS * 32)
        if (base + i < n) atomicAdd(&h[(key[base + i] >> shift) & 255u], 1u);
    __syncthreads();
    hist[(uint64_t)threadIdx.x * tiles + blockIdx.x] = h[threadIdx.x];
}

__global__ void __launch_bounds__(RS_WARPS * 32) k_radix_scatter(const uint32_t *key, const uint32_t *val, uint32_t *key_out, uint32_t *val_out, uint64_t n, int shift,
                                                                 const uint32_t *hist_scanned, uint32_t tiles)
{
    __shared__ uint32_t cnt[RS_WARPS][256];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_WARPS * 32) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * RS_TILE + (uint64_t)w * 32 * RS_ROUNDS;
    uint32_t k[RS_ROUNDS], v[RS_ROUNDS];
    bool in[RS_ROUNDS];
#pragma unroll
    for (int q = 0; q < RS_ROUNDS; q++) {
        const uint64_t i = base + q * 32 + lane;
        in[q] = i < n;
        k[q] = in[q] ? key[i] : 0xffffffffu, v[q] = in[q] ? val[i] : 0u;
        if (in[q]) atomicAdd(&cnt[w][(k[q] >> shift) & 255u], 1u);
    }
    __syncthreads();
    { // digit d (one thread each): where warp w's first pair with this digit goes
        const int d = threadIdx.x;
        uint32_t run = hist_scanned[(uint64_t)d * tiles + blockIdx.x];
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ww++) {
            const uint32_t c = cnt[ww][d];
            cnt[ww][d] = run, run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < RS_ROUNDS; q++) {
        const uint32_t d = (k[q] >> shift) & 255u;
        const unsigned act = __ballot_sync(0xffffffffu, in[q]);
        if (in[q]) {
            const unsigned same = __match_any_sync(act, d);
            const uint32_t at = cnt[w][d] + __popc(same & ((1u << lane) - 1u));
            key_out[at] = k[q], val_out[at] = v[q];
            __syncwarp(same);
            if (lane == __ffs(same) - 1) cnt[w][d] += __popc(same);
        }
        __syncwarp(); // the next round may meet a digit another group of lanes has just advanced
    }
}

// sorts n pairs by the low `bits` bits of the key; the result ends in (key, val) (the buffers are swapped an even number of times or copied back)
cudaError_t radix_sort_pairs(uint32_t *key, uint32_t *val, uint64_t n, int bits, scratch &sc, cudaStream_t s)
{
    if (n <= 1 || bits <= 0) return cudaSuccess;
    const uint32_t tiles = blocks_for(n, RS_TILE);
    uint32_t *key2 = nullptr, *val2 = nullptr, *hist = nullptr;
    TRY(sc.get(&key2, n));
    TRY(sc.get(&val2, n));
    TRY(sc.get(&hist, (uint64_t)256 * tiles + 1));
    uint32_t *ka = key, *va = val, *kb = key2, *vb = val2;
    for (int shift = 0; shift < bits; shift += 8) {
        k_radix_hist<<<tiles, RS_WARPS * 32, 0, s>>>(ka, n, shift, hist, tiles);
        TRY(cudaGetLastError());
        TRY(scan_exclusive(hist, hist, (uint64_t)256 * tiles, sc, s));
        k_radix_scatter<<<tiles, RS_WARPS * 32, 0, s>>>(ka, va, kb, vb, n, shift, hist, tiles);
        TRY(cudaGetLastError());
        std::swap(ka, kb), std::swap(va, vb);
    }
    if (ka != key) {
        TRY(cudaMemcpyAsync(key, ka, n * 4, cudaMemcpyDeviceToDevice, s));
        TRY(cudaMemcpyAsync(val, va, n * 4, cudaMemcpyDeviceToDevice, s));
    }
    return cudaSuccess;
}

// ---- passes shared by the chains ---------------------------------------------------------------------------------------------------------------
// flags[0]: a row id below its predecessor or >= rows, or a column id >= cols
__global__ void k_check_coo(const uint32_t *row, const uint32_t *col, uint64_t nnz, uint32_t rows, uint32_t cols, uint32_t *flags)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (uint64_t)gridDim.x * blockDim.x)
        if (row[i] >= rows || col[i] >= cols || (i && row[i] < row[i - 1])) flags[0] = 1;
}

// CSR offsets of a row-major COO: rowptr[r] = first stream position whose row id is >= r, r in [0, rows]
// (get_nnz_of_each_row_in_spec_range, data_transform_common.cc:7-50, as offsets)
__global__ void k_rowptr(const uint32_t *row, uint32_t nnz, uint32_t rows, uint32_t *rowptr)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= rows; r += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = nnz;
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (row[mid] < r) lo = mid + 1;
            else hi = mid;
        }
        rowptr[r] = lo;
    }
}

__global__ void k_lengths(const uint32_t *rowptr, uint32_t rows, uint32_t *len)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (uint64_t)gridDim.x * blockDim.x) len[r] = rowptr[r + 1] - rowptr[r];
}

// sort keys: longest row first
__global__ void k_sort_keys(const uint32_t *len, uint32_t rows, uint32_t max_len, uint32_t *key, uint32_t *val)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (uint64_t)gridDim.x * blockDim.x) key[r] = max_len - len[r], val[r] = (uint32_t)r;
}

__global__ void k_iota(uint32_t *out, uint64_t n, uint32_t step)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = (uint32_t)i * step;
}

// padded length of every kept row, in the chain's row order (modify_col_indices_by_col_pad_in_sub_matrix.cc: rows whose length is not a
// multiple of m grow to the next one; empty rows stay empty)
__global__ void k_padded_len(const uint32_t *len, const uint32_t *order, uint32_t rows_kept, uint32_t m, uint32_t *plen)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows_kept; r += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t l = len[order ? order[r] : r];
        plen[r] = (m > 1 && l % m) ? (l / m + 1) * m : l;
    }
}

// one warp per kept row: the row's entries move to their place in the (sorted, padded) stream (reorder_{val,col,row}_by_index.cc, then the
// pad entries of modify_*_by_col_pad_in_sub_matrix.cc: column = the row's last column, value 0, the same row)
template <typename ColT, typename ValT>
__global__ void __launch_bounds__(256) k_scatter_rows(const uint32_t *col, const float *val, const uint32_t *rowptr, const uint32_t *order, const uint32_t *rowptr_out, uint32_t rows_kept,
                                                      ColT *col_out, ValT *val_out, uint32_t *row_out)
{
    const int lane = threadIdx.x & 31;
    for (uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows_kept; r += ((uint64_t)gridDim.x * blockDim.x) >> 5) {
        const uint32_t o = order ? order[r] : (uint32_t)r;
        const uint32_t src = rowptr[o], l = rowptr[o + 1] - src, dst = rowptr_out[r], pl = rowptr_out[r + 1] - dst;
        const uint32_t last = l ? col[src + l - 1] : 0u;
        for (uint32_t i = lane; i < pl; i += 32) {
            col_out[dst + i] = (ColT)(i < l ? col[src + i] : last);
            const float v = i < l ? val[src + i] : 0.0f;
            if (sizeof(ValT) == 2) ((__half *)val_out)[dst + i] = __float2half_rn(v);
            else ((float *)val_out)[dst + i] = v;
            row_out[dst + i] = (uint32_t)r;
        }
    }
}

// ---- nnz direction ----------------------------------------------------------------------------------------------------------------------------
// modify_{col_indices,vals,row_indices}_by_nnz_pad.cc: the stream grows to a multiple of the block with copies of the last column / row and value 0
template <typename ColT, typename ValT>
__global__ void k_pad_stream(const uint32_t *row, const uint32_t *col, const float *val, uint32_t nnz, uint32_t nnz_pad, ColT *col_out, ValT *val_out, uint32_t *row_out)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz_pad; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t j = i < nnz ? (uint32_t)i : nnz - 1;
        col_out[i] = (ColT)col[j], row_out[i] = row[j];
        const float v = i < nnz ? val[i] : 0.0f;
        if (sizeof(ValT) == 2) ((__half *)val_out)[i] = __float2half_rn(v);
        else ((float *)val_out)[i] = v;
    }
}

// One warp per BMT of 32 non-zeros (row ids of the PADDED stream).  Pass 1: first_row_indices (get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_
// direction.cc), thread_bit_map (thread_bit_map.cc: bit j = non-zero j of the BMT starts a row, LSB first; the stream's first non-zero always does),
// the BMT's segment count (segment_ptr.cc: one plus the row changes inside the BMT) and its gap flag (segment_empty_flag.cc: two neighbours inside
// the BMT more than one row apart).  Pass 2, after the scan of the counts: segment_empty_row_indices (segment_empty_row_indices.cc: 0, then the row
// offset of every row change from the BMT's first row).
// parent: thread_bit_map.cc's parent_flag loop (thread_bit_map_operator at WARP level): the first non-zero of every `parent`-th BMT starts a segment too (0 = off)
__global__ void __launch_bounds__(256) k_bmt_pass(const uint32_t *row_pad, uint32_t n_bmt, uint32_t *first_row, uint64_t *bit_map, uint32_t *seg_count, uint8_t *gap_flag,
                                                  const uint32_t *seg_ptr, uint32_t *seg_rows, uint32_t parent)
{
    const int lane = threadIdx.x & 31;
    for (uint64_t b = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_bmt; b += ((uint64_t)gridDim.x * blockDim.x) >> 5) {
        const uint64_t i = b * 32 + lane;
        const uint32_t r = row_pad[i], prev = i ? row_pad[i - 1] : 0u;
        const bool start = i == 0 || r != prev;
        const unsigned starts = __ballot_sync(0xffffffffu, start);
        const unsigned inner = starts & ~1u; // row changes after the BMT's first non-zero
        const unsigned gaps = __ballot_sync(0xffffffffu, lane > 0 && r - prev > 1);
        const uint32_t r0 = __shfl_sync(0xffffffffu, r, 0);
        if (!seg_ptr) {
            if (lane == 0) first_row[b] = r0, bit_map[b] = (uint64_t)(starts | ((parent && b % parent == 0) ? 1u : 0u)), seg_count[b] = 1u + __popc(inner), gap_flag[b] = gaps ? 1 : 0;
        } else {
            const uint32_t at = seg_ptr[b];
            if (lane == 0) seg_rows[at] = 0;
            else if (start) seg_rows[at + 1 + __popc(inner & ((1u << lane) - 1u))] = r - r0;
        }
    }
}

// segment_offset.cc: a BMT whose bit map is not zero gets the number of all-zero BMTs between it and the next non-zero one; the run that reaches
// the end of the array is never written back (kept: the reference's loop ends without flushing it)
template <typename W> __global__ void k_segment_offset(const W *bit_map, uint32_t n_bmt, uint32_t *out)
{
    for (uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_bmt; b += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t v = 0;
        if (b == 0 || bit_map[b] != 0) {
            uint64_t j = b + 1;
            while (j < n_bmt && bit_map[j] == 0) j++;
            if (j < n_bmt) v = (uint32_t)(j - b - 1);
        }
        out[b] = v;
    }
}

// get_begin_BMTs_after_merge_thread.cc: [0, m, 2 m, ..., n]
__global__ void k_group_starts(uint32_t *out, uint32_t n_out, uint32_t m, uint32_t n)
{
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_out; j += (uint64_t)gridDim.x * blockDim.x) out[j] = (uint32_t)min((uint64_t)j * m, (uint64_t)n);
}

// segment_empty_flag.cc: `size` consecutive BMT flags per word, the first in the LSB
__global__ void k_pack_flags(const uint8_t *flag, uint32_t n_bmt, uint32_t size, uint64_t *out, uint32_t n_words)
{
    for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t word = 0;
        for (uint32_t j = 0; j < size && w * size + j < n_bmt; j++) word |= (uint64_t)flag[w * size + j] << j;
        out[w] = word;
    }
}

// ---- column direction (rows padded to whole BMTs of `c` non-zeros) ------------------------------------------------------------------------------
// get_begin_{rows,nzs}_of_BMT_after_fixed_blocking_in_col_direction.cc: BMT b is non-zeros [c b, c b + c) and lies in the row of its first one;
// get_begin_nzs_relative_to_parent_after_merge_thread.cc: the same offsets counted from the first BMT of the group of `m`
__global__ void k_col_bmts(const uint32_t *row_pad, uint32_t n_bmt, uint32_t c, uint32_t m, uint32_t *bmt_row, uint32_t *first_nz, uint32_t *nz_rel)
{
    for (uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; b <= n_bmt; b += (uint64_t)gridDim.x * blockDim.x) {
        first_nz[b] = (uint32_t)b * c;
        if (b < n_bmt) {
            bmt_row[b] = row_pad[b * c];
            if (nz_rel) nz_rel[b] = (uint32_t)(b % m) * c;
        }
    }
}
// get_begin_rows_after_merge_thread.cc (sample_with_last): every m-th entry while another follows, then the last one
__global__ void k_sample_with_last(const uint32_t *in, uint32_t n_in, uint32_t m, uint32_t *out, uint32_t n_out)
{
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_out; j += (uint64_t)gridDim.x * blockDim.x) out[j] = j + 1 < n_out ? in[j * m] : in[n_in - 1];
}
// parent_bit_map_of_thread.cc: a BMT starts a segment when its row differs from the BMT before it, and every m-th BMT does.  WARP level: m bits per
// u64 word, first BMT in the LSB; TBLOCK level: one 0 / 1 per BMT
__global__ void k_parent_bits(const uint32_t *bmt_row, uint32_t n_bmt, uint32_t m, uint64_t *words, uint32_t n_words, uint32_t *flags)
{
    const uint64_t n = words ? n_words : n_bmt;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (words) {
            uint64_t w = 0;
            for (uint32_t j = 0; j < m && i * m + j < n_bmt; j++) {
                const uint64_t b = i * m + j;
                if (j == 0 || bmt_row[b] != bmt_row[b - 1]) w |= 1ull << j;
            }
            words[i] = w;
        } else flags[i] = (i == 0 || i % m == 0 || bmt_row[i] != bmt_row[i - 1]) ? 1u : 0u;
    }
}

cudaError_t take(transform_out &out, const char *pos, const char *name, void *ptr, uint64_t len, int dtype, uint64_t max_value, uint64_t align, bool iota = false, bool linear = false,
                 uint64_t step = 0)
{
    built_array a;
    a.pos = pos, a.name = name, a.ptr = ptr, a.len = len, a.dtype = dtype, a.max_value = max_value, a.align = align, a.iota = iota, a.linear = linear, a.linear_step = step;
    out.arrays.push_back(a);
    return cudaSuccess;
}

template <typename T> cudaError_t alloc_out(T **p, uint64_t n, cudaStream_t s) { return pool_alloc((void **)p, std::max<uint64_t>(n, 1) * sizeof(T) + 16, s); }

cudaError_t check_input(const coo_dev &m, scratch &sc, cudaStream_t s, transform_out &out, uint32_t *max_col)
{
    if (!m.row || !m.col || !m.val || m.nnz == 0 || m.rows == 0 || m.cols == 0 || m.nnz > 0xfffffff0ull || m.rows > 0xfffffff0ull || m.cols > 0xffffffffull) {
        out.error = "device transform: empty matrix or more than 2^32 - 16 non-zeros / rows";
        return cudaErrorInvalidValue;
    }
    uint32_t *flag = nullptr, bad = 0;
    TRY(sc.get(&flag, 1));
    TRY(cudaMemsetAsync(flag, 0, 4, s));
    k_check_coo<<<blocks_for(m.nnz, 256 * 8), 256, 0, s>>>(m.row, m.col, m.nnz, (uint32_t)m.rows, (uint32_t)m.cols, flag);
    TRY(cudaGetLastError());
    TRY(cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, s));
    TRY(reduce_u32<op_max>(m.col, m.nnz, max_col, sc, s));
    if (bad) {
        out.error = "device transform: the COO stream must be in row-major order with ids inside the matrix";
        return cudaErrorInvalidValue;
    }
    return cudaSuccess;
}

bool over_pad_bound(uint64_t after, uint64_t before, double bound, const char *who, transform_out &out)
{
    // the reference asserts when nnz_after / nnz_before >= PADDING_RATE_UP_BOUND (the count only grows, so the final ratio decides)
    if (after == before || (double)after / (double)before < bound) return false;
    char buf[200];
    snprintf(buf, sizeof buf, "%s: current padding rate is %f, higher than %f", who, (double)after / (double)before, bound);
    out.error = buf, out.refused = 1;
    return true;
}

} // namespace

namespace {
__global__ void k_set_u32(uint32_t *p, uint32_t v) { *p = v; }
__global__ void k_flag_nonzero(const uint32_t *in, uint32_t n, uint32_t *out)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = in[i] != 0;
}
__global__ void k_not_iota(const uint32_t *in, uint64_t n, uint32_t *flag)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        if (in[i] != (uint32_t)i) *flag = 1;
}
__global__ void k_first_rows(const uint32_t *row_pad, uint32_t n_bmt, uint32_t rows, uint32_t *first_row)
{
    for (uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; b <= n_bmt; b += (uint64_t)gridDim.x * blockDim.x) first_row[b] = b < n_bmt ? row_pad[b * 32] : rows;
}
int bit_width(uint32_t v)
{
    int b = 0;
    while (v) b++, v >>= 1;
    return b;
}
} // namespace

// exported for the other plan-time builders (kernels_merge.cu): out[i] = in[0] + ... + in[i - 1], i in [0, n]
cudaError_t scan_exclusive_u32(const uint32_t *in, uint32_t *out, uint64_t n, cudaStream_t s)
{
    scratch sc(s);
    return scan_exclusive(in, out, n, sc, s);
}
cudaError_t rowptr_from_sorted_rows(const uint32_t *row, uint32_t nnz, uint32_t rows, uint32_t *rowptr, cudaStream_t s)
{
    k_rowptr<<<blocks_for((uint64_t)rows + 1, 256), 256, 0, s>>>(row, nnz, rows, rowptr);
    return cudaGetLastError();
}

cudaError_t transform_row_chain(const coo_dev &m, bool sort, uint32_t S, bool half, const char *level, double pad_bound, cudaStream_t s, transform_out &out)
{
    scratch sc(s);
    uint32_t max_col = 0;
    TRY(check_input(m, sc, s, out, &max_col));
    if (S == 1) { // modify_col_indices_by_col_pad_in_sub_matrix.cc asserts multiple_of_each_row_size >= 2
        out.error = "modify_by_col_pad_in_sub_matrix: multiple_of_each_row_size must be >= 2";
        return cudaErrorInvalidValue;
    }
    const uint32_t rows = (uint32_t)m.rows, nnz = (uint32_t)m.nnz;
    uint32_t *rowptr = nullptr, *len = nullptr, *plen = nullptr;
    TRY(sc.get(&rowptr, (uint64_t)rows + 1));
    TRY(sc.get(&len, rows));
    TRY(sc.get(&plen, (uint64_t)rows + 1));
    k_rowptr<<<blocks_for((uint64_t)rows + 1, 256), 256, 0, s>>>(m.row, nnz, rows, rowptr);
    k_lengths<<<blocks_for(rows, 256 * 4), 256, 0, s>>>(rowptr, rows, len);
    TRY(cudaGetLastError());

    uint32_t *order = nullptr, rows_kept = rows;
    if (sort) { // sort_operator.cc:70-108
        uint32_t max_len = 0, nonempty = 0;
        TRY(reduce_u32<op_max>(len, rows, &max_len, sc, s));
        uint32_t *key = nullptr;
        TRY(sc.get(&key, rows));
        TRY(alloc_out(&order, rows, s));
        take(out, "GLOBAL_META", "original_nz_row_indices", order, rows, GS_U32, rows - 1, 1);
        k_sort_keys<<<blocks_for(rows, 256 * 4), 256, 0, s>>>(len, rows, max_len, key, order);
        TRY(cudaGetLastError());
        TRY(radix_sort_pairs(key, order, rows, bit_width(max_len), sc, s));
        // remove_empty_row_in_end_of_sub_matrix.cc: the sorted stream ends with its last non-empty row; rows behind it leave the row level
        k_flag_nonzero<<<blocks_for(rows, 256 * 4), 256, 0, s>>>(len, rows, plen);
        TRY(cudaGetLastError());
        TRY(reduce_u32<op_sum>(plen, rows, &nonempty, sc, s));
        rows_kept = nonempty;
        out.has_perm = true;
    }
    out.rows_kept = rows_kept;

    uint32_t *first_nz = nullptr, *first_row = nullptr;
    TRY(alloc_out(&first_nz, (uint64_t)rows_kept + 1, s));
    take(out, level, "first_nz_indices", first_nz, (uint64_t)rows_kept + 1, GS_U32, 0, 0);
    TRY(alloc_out(&first_row, (uint64_t)rows_kept + 1, s));
    take(out, level, "first_row_indices", first_row, (uint64_t)rows_kept + 1, GS_U32, rows_kept, 1, true, true, 1);
    k_padded_len<<<blocks_for(rows_kept, 256 * 4), 256, 0, s>>>(len, order, rows_kept, S, plen);
    k_iota<<<blocks_for((uint64_t)rows_kept + 1, 256 * 4), 256, 0, s>>>(first_row, (uint64_t)rows_kept + 1, 1);
    TRY(cudaGetLastError());
    uint32_t plen_max = 0, plen_gcd = 0, nnz_pad = 0;
    TRY(reduce_u32<op_max>(plen, rows_kept, &plen_max, sc, s));
    TRY(reduce_u32<op_gcd>(plen, rows_kept, &plen_gcd, sc, s));
    TRY(scan_exclusive(plen, first_nz, rows_kept, sc, s));
    TRY(cudaMemcpyAsync(&nnz_pad, first_nz + rows_kept, 4, cudaMemcpyDeviceToHost, s));
    TRY(cudaStreamSynchronize(s));
    if (over_pad_bound(nnz_pad, nnz, pad_bound, "modify_col_indices_by_col_pad_in_sub_matrix", out)) return cudaErrorInvalidValue;
    {
        built_array &fn = out.arrays[out.arrays.size() - 2];
        fn.max_value = nnz_pad, fn.align = plen_gcd;
        fn.linear = (uint64_t)plen_max * rows_kept == nnz_pad, fn.linear_step = plen_max; // every row as long as the longest
        fn.iota = fn.linear && plen_max == 1;
    }
    out.stored_nnz = nnz_pad;

    const bool c16 = max_col <= 65535u;
    void *col_out = nullptr, *val_out = nullptr;
    uint32_t *row_out = nullptr;
    TRY(pool_alloc(&col_out, (uint64_t)nnz_pad * (c16 ? 2 : 4) + 16, s));
    take(out, "GLOBAL_META", "nz_col_indices", col_out, nnz_pad, c16 ? GS_U16 : GS_U32, max_col, 0);
    TRY(pool_alloc(&val_out, (uint64_t)nnz_pad * (half ? 2 : 4) + 16, s));
    take(out, "GLOBAL_META", "nz_vals", val_out, nnz_pad, half ? GS_F16 : GS_F32, 0, 0);
    TRY(alloc_out(&row_out, nnz_pad, s));
    take(out, "GLOBAL_META", "nz_row_indices", row_out, nnz_pad, GS_U32, rows_kept ? rows_kept - 1 : 0, 0);
    const unsigned grid = std::min(blocks_for((uint64_t)rows_kept * 32, 256), 148u * 64u);
    if (c16 && half) k_scatter_rows<uint16_t, __half><<<grid, 256, 0, s>>>(m.col, m.val, rowptr, order, first_nz, rows_kept, (uint16_t *)col_out, (__half *)val_out, row_out);
    else if (c16) k_scatter_rows<uint16_t, float><<<grid, 256, 0, s>>>(m.col, m.val, rowptr, order, first_nz, rows_kept, (uint16_t *)col_out, (float *)val_out, row_out);
    else if (half) k_scatter_rows<uint32_t, __half><<<grid, 256, 0, s>>>(m.col, m.val, rowptr, order, first_nz, rows_kept, (uint32_t *)col_out, (__half *)val_out, row_out);
    else k_scatter_rows<uint32_t, float><<<grid, 256, 0, s>>>(m.col, m.val, rowptr, order, first_nz, rows_kept, (uint32_t *)col_out, (float *)val_out, row_out);
    TRY(cudaGetLastError());
    return cudaStreamSynchronize(s);
}

cudaError_t transform_nnz_chain(const coo_dev &m, uint32_t size, bool warp_level, bool half, double pad_bound, cudaStream_t s, transform_out &out)
{
    scratch sc(s);
    uint32_t max_col = 0;
    TRY(check_input(m, sc, s, out, &max_col));
    if (size < 1 || size > 64) {
        out.error = "segment_empty_flag packs at most 64 BMTs per word";
        return cudaErrorInvalidValue;
    }
    const uint32_t rows = (uint32_t)m.rows, nnz = (uint32_t)m.nnz;
    const uint64_t padded = nnz % 32 ? ((uint64_t)nnz / 32 + 1) * 32 : nnz;
    if (padded > 0xfffffff0ull) {
        out.error = "device transform: the padded stream does not fit 32-bit offsets";
        return cudaErrorInvalidValue;
    }
    if (over_pad_bound(padded, nnz, pad_bound, "modify_col_indices_by_nnz_pad", out)) return cudaErrorInvalidValue;
    const uint32_t nnz_pad = (uint32_t)padded, n_bmt = nnz_pad / 32;
    out.stored_nnz = nnz_pad, out.rows_kept = rows;

    const bool c16 = max_col <= 65535u;
    void *col_out = nullptr, *val_out = nullptr;
    uint32_t *row_out = nullptr;
    TRY(pool_alloc(&col_out, (uint64_t)nnz_pad * (c16 ? 2 : 4) + 16, s));
    take(out, "GLOBAL_META", "nz_col_indices", col_out, nnz_pad, c16 ? GS_U16 : GS_U32, max_col, 0);
    TRY(pool_alloc(&val_out, (uint64_t)nnz_pad * (half ? 2 : 4) + 16, s));
    take(out, "GLOBAL_META", "nz_vals", val_out, nnz_pad, half ? GS_F16 : GS_F32, 0, 0);
    TRY(alloc_out(&row_out, nnz_pad, s));
    take(out, "GLOBAL_META", "nz_row_indices", row_out, nnz_pad, GS_U32, rows - 1, 0);
    const unsigned g1 = std::min(blocks_for(nnz_pad, 256 * 4), 148u * 64u);
    if (c16 && half) k_pad_stream<uint16_t, __half><<<g1, 256, 0, s>>>(m.row, m.col, m.val, nnz, nnz_pad, (uint16_t *)col_out, (__half *)val_out, row_out);
    else if (c16) k_pad_stream<uint16_t, float><<<g1, 256, 0, s>>>(m.row, m.col, m.val, nnz, nnz_pad, (uint16_t *)col_out, (float *)val_out, row_out);
    else if (half) k_pad_stream<uint32_t, __half><<<g1, 256, 0, s>>>(m.row, m.col, m.val, nnz, nnz_pad, (uint32_t *)col_out, (__half *)val_out, row_out);
    else k_pad_stream<uint32_t, float><<<g1, 256, 0, s>>>(m.row, m.col, m.val, nnz, nnz_pad, (uint32_t *)col_out, (float *)val_out, row_out);
    TRY(cudaGetLastError());

    uint32_t *first_row = nullptr, *first_nz = nullptr, *seg_ptr = nullptr, *seg_count = nullptr, *seg_rows = nullptr, *flag = nullptr;
    uint64_t *bit_map = nullptr, *gap_words = nullptr;
    uint8_t *gap = nullptr;
    TRY(alloc_out(&first_row, (uint64_t)n_bmt + 1, s));
    take(out, "THREAD_META", "first_row_indices", first_row, (uint64_t)n_bmt + 1, GS_U32, rows, 0);
    TRY(alloc_out(&first_nz, (uint64_t)n_bmt + 1, s));
    take(out, "THREAD_META", "first_nz_indices", first_nz, (uint64_t)n_bmt + 1, GS_U32, nnz_pad, 32, false, true, 32);
    TRY(alloc_out(&bit_map, n_bmt, s));
    take(out, "THREAD_META", "thread_bit_map", bit_map, n_bmt, GS_U64, 0, 0);
    TRY(alloc_out(&seg_ptr, (uint64_t)n_bmt + 1, s));
    take(out, "THREAD_META", "segment_ptr", seg_ptr, n_bmt, GS_U32, 0, 0);
    TRY(sc.get(&seg_count, n_bmt));
    TRY(sc.get(&gap, n_bmt));
    TRY(sc.get(&flag, 1));
    const unsigned g2 = std::min(blocks_for((uint64_t)n_bmt * 32, 256), 148u * 64u);
    k_first_rows<<<blocks_for((uint64_t)n_bmt + 1, 256 * 4), 256, 0, s>>>(row_out, n_bmt, rows, first_row);
    k_iota<<<blocks_for((uint64_t)n_bmt + 1, 256 * 4), 256, 0, s>>>(first_nz, (uint64_t)n_bmt + 1, 32);
    k_bmt_pass<<<g2, 256, 0, s>>>(row_out, n_bmt, first_row, bit_map, seg_count, gap, nullptr, nullptr, warp_level ? size : 0u);
    TRY(cudaGetLastError());
    TRY(scan_exclusive(seg_count, seg_ptr, n_bmt, sc, s));
    uint32_t n_seg_rows = 0, first_row_gcd = 0, not_iota = 0;
    TRY(cudaMemsetAsync(flag, 0, 4, s));
    k_not_iota<<<blocks_for((uint64_t)n_bmt + 1, 256 * 4), 256, 0, s>>>(first_row, (uint64_t)n_bmt + 1, flag);
    TRY(cudaMemcpyAsync(&not_iota, flag, 4, cudaMemcpyDeviceToHost, s));
    TRY(cudaMemcpyAsync(&n_seg_rows, seg_ptr + n_bmt, 4, cudaMemcpyDeviceToHost, s));
    TRY(reduce_u32<op_gcd>(first_row, (uint64_t)n_bmt + 1, &first_row_gcd, sc, s)); // (synchronizes)
    {
        built_array &fr = out.arrays[out.arrays.size() - 4];
        fr.align = first_row_gcd, fr.iota = !not_iota, fr.linear = fr.iota, fr.linear_step = 1;
        out.arrays.back().max_value = n_seg_rows;
    }
    TRY(alloc_out(&seg_rows, n_seg_rows, s));
    take(out, "THREAD_META", "segment_empty_row_indices", seg_rows, n_seg_rows, GS_U32, 0, 0);
    const uint32_t n_words = (n_bmt + size - 1) / size;
    TRY(alloc_out(&gap_words, n_words, s));
    take(out, "THREAD_META", "segment_empty_flag", gap_words, n_words, GS_U64, 0, 0);
    k_bmt_pass<<<g2, 256, 0, s>>>(row_out, n_bmt, nullptr, nullptr, nullptr, nullptr, seg_ptr, seg_rows, 0u);
    k_pack_flags<<<blocks_for(n_words, 256), 256, 0, s>>>(gap, n_bmt, size, gap_words, n_words);
    TRY(cudaGetLastError());
    if (warp_level) { // thread_bit_map_operator(WARP_META) + warp_segment_reduce_operator: segment_offset and the BMWs of `size` BMTs
        uint32_t *seg_off = nullptr, *first_bmt = nullptr;
        TRY(alloc_out(&seg_off, n_bmt, s));
        take(out, "THREAD_META", "segment_offset", seg_off, n_bmt, GS_U32, 0, 0);
        const uint32_t n_w = (n_bmt + size - 1) / size + 1;
        TRY(alloc_out(&first_bmt, n_w, s));
        const uint32_t g = n_bmt % size ? 1u : size;
        take(out, "WARP_META", "first_BMT_indices", first_bmt, n_w, GS_U32, n_bmt, n_bmt % size ? std::__gcd(size, n_bmt) : size, size == 1, n_bmt % size == 0, size);
        (void)g;
        k_segment_offset<uint64_t><<<blocks_for(n_bmt, 256), 256, 0, s>>>(bit_map, n_bmt, seg_off);
        k_group_starts<<<blocks_for(n_w, 256), 256, 0, s>>>(first_bmt, n_w, size, n_bmt);
        TRY(cudaGetLastError());
    }
    return cudaStreamSynchronize(s);
}

} // namespace gsd

namespace gsd {

// fixed_interval_col_direction_thread_blocking_operator(64, pad to the column block) + thread_total_reduce_operator, then
//   warp_level:  warp_bit_map_operator(C, relative nz, relative row) -- BMWs of `group` BMTs: WARP first_{row,nz,BMT}_indices, WARP bit_map_of_thread, THREAD
//                first_nz_indices_relative_to_BMW, WARP first_row_indices_relative_to_BMW
//   !warp_level: tblock_thread_bit_map_operator(C, group) -- BMTBs of `group` BMTs: TBLOCK first_BMT_indices, THREAD bit_map_of_thread, THREAD segment_offset
cudaError_t transform_col_chain(const coo_dev &m, uint32_t col_block, uint32_t group, bool warp_level, bool half, double pad_bound, cudaStream_t s, transform_out &out)
{
    TRY(transform_row_chain(m, false, col_block, half, "ROW_TMP", pad_bound, s, out)); // (the padding refusal comes first, as on the host)
    if (group < 1 || (warp_level && group > 64)) { // parent_bit_map_of_thread.cc packs VECTOR_WIDTH BMTs per u64 word
        out.error = "parent_bit_map_of_thread: VECTOR_WIDTH must be in 1..64";
        return cudaErrorInvalidValue;
    }
    // the row level of the padded stream is scaffolding here: the chain registers BMT-level arrays only
    uint32_t *row_pad = nullptr;
    for (size_t i = 0; i < out.arrays.size();) {
        if (out.arrays[i].pos == "ROW_TMP") {
            cudaFreeAsync(out.arrays[i].ptr, s);
            out.arrays.erase(out.arrays.begin() + i);
        } else {
            if (out.arrays[i].name == "nz_row_indices") row_pad = (uint32_t *)out.arrays[i].ptr;
            i++;
        }
    }
    const uint32_t nnz_pad = (uint32_t)out.stored_nnz, n_bmt = nnz_pad / col_block;
    if (!row_pad || n_bmt == 0 || nnz_pad % col_block) return cudaErrorInvalidValue;
    uint32_t *bmt_row = nullptr, *first_nz = nullptr, *nz_rel = nullptr, *first_bmt = nullptr;
    TRY(alloc_out(&bmt_row, n_bmt, s));
    take(out, "THREAD_META", "first_row_indices_without_ending", bmt_row, n_bmt, GS_U32, m.rows - 1, 0);
    // (the warp_bitmap chain registers the BMT offsets relative to their BMW only: the absolute ones are scratch there, as in the host's upload)
    scratch sc(s);
    if (warp_level) TRY(sc.get(&first_nz, (uint64_t)n_bmt + 1));
    else {
        TRY(alloc_out(&first_nz, (uint64_t)n_bmt + 1, s));
        take(out, "THREAD_META", "first_nz_indices", first_nz, (uint64_t)n_bmt + 1, GS_U32, nnz_pad, col_block, false, true, col_block);
    }
    if (warp_level) {
        TRY(alloc_out(&nz_rel, n_bmt, s));
        take(out, "THREAD_META", "first_nz_indices_relative_to_BMW", nz_rel, n_bmt, GS_U32, (uint64_t)(std::min(group, n_bmt) - 1) * col_block, group > 1 ? col_block : 0);
    }
    k_col_bmts<<<blocks_for((uint64_t)n_bmt + 1, 256 * 4), 256, 0, s>>>(row_pad, n_bmt, col_block, group, bmt_row, first_nz, nz_rel);
    TRY(cudaGetLastError());
    const uint32_t n_g = (n_bmt + group - 1) / group + 1; // [0, group, ..., n_bmt]
    TRY(alloc_out(&first_bmt, n_g, s));
    take(out, warp_level ? "WARP_META" : "TBLOCK_META", "first_BMT_indices", first_bmt, n_g, GS_U32, n_bmt, n_bmt % group ? std::__gcd(group, n_bmt) : group, group == 1, n_bmt % group == 0, group);
    k_group_starts<<<blocks_for(n_g, 256), 256, 0, s>>>(first_bmt, n_g, group, n_bmt);
    TRY(cudaGetLastError());
    if (warp_level) {
        uint32_t *w_row = nullptr, *w_nz = nullptr, *row_rel = nullptr;
        uint64_t *bits = nullptr;
        const uint32_t n_wr = (n_bmt - 1 + group - 1) / group + 1; // sample_with_last over n_bmt entries: one entry fewer than the offsets when (n_bmt - 1) % group == 0
        TRY(alloc_out(&w_row, n_wr, s));
        take(out, "WARP_META", "first_row_indices", w_row, n_wr, GS_U32, m.rows - 1, 0);
        TRY(alloc_out(&w_nz, n_g, s));
        take(out, "WARP_META", "first_nz_indices", w_nz, n_g, GS_U32, nnz_pad, 0);
        const uint32_t n_words = (n_bmt + group - 1) / group;
        TRY(alloc_out(&bits, n_words, s));
        take(out, "WARP_META", "bit_map_of_thread", bits, n_words, GS_U64, 0, 0);
        k_sample_with_last<<<blocks_for(n_wr, 256), 256, 0, s>>>(bmt_row, n_bmt, group, w_row, n_wr);
        k_sample_with_last<<<blocks_for(n_g, 256), 256, 0, s>>>(first_nz, n_bmt + 1, group, w_nz, n_g);
        k_parent_bits<<<blocks_for(n_words, 256), 256, 0, s>>>(bmt_row, n_bmt, group, bits, n_words, nullptr);
        TRY(cudaGetLastError());
        (void)row_rel;
    } else {
        uint32_t *flags = nullptr, *seg_off = nullptr;
        TRY(alloc_out(&flags, n_bmt, s));
        take(out, "THREAD_META", "bit_map_of_thread", flags, n_bmt, GS_U32, 1, 0);
        TRY(alloc_out(&seg_off, n_bmt, s));
        take(out, "THREAD_META", "segment_offset", seg_off, n_bmt, GS_U32, 0, 0);
        k_parent_bits<<<blocks_for(n_bmt, 256), 256, 0, s>>>(bmt_row, n_bmt, group, nullptr, 0, flags);
        k_segment_offset<uint32_t><<<blocks_for(n_bmt, 256), 256, 0, s>>>(flags, n_bmt, seg_off);
        TRY(cudaGetLastError());
    }
    return cudaStreamSynchronize(s);
}

} // namespace gsd

namespace gsd {
namespace {
// merge path (get_begin_{rows,nzs}_of_level_after_merge_path.cc; csrc/host/steps.cc: merge_path_walk): the path visits the non-empty rows, row j (the
// j-th non-empty one) ends at path position sum(len of rows 0..j) + j; part p starts at position p * work, in the first row whose end lies
// beyond it, at non-zero position - j.  With incl[r] = sum over rows <= r of (len + 1 where len > 0) the end of a non-empty row r is incl[r] - 1,
// incl only grows at non-empty rows, so the row of part p is the first r with incl[r] >= p * work + 2 -- one binary search per part.
__global__ void k_mp_path_terms(const uint32_t *rowptr, uint32_t n_rows, uint32_t *term, uint32_t *nonempty)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t l = rowptr[r + 1] - rowptr[r];
        term[r] = l ? l + 1 : 0, nonempty[r] = l != 0;
    }
}
__global__ void k_mp_parts(const uint32_t *excl /* n_rows + 1: excl[r + 1] = incl[r] */, const uint32_t *ordinal, uint32_t n_rows, uint32_t work, uint32_t n_parts, uint32_t nnz,
                           uint32_t *part_row, uint32_t *part_nz)
{
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p <= n_parts; p += (uint64_t)gridDim.x * blockDim.x) {
        if (p == n_parts) {
            part_nz[p] = nnz; // the sentinel get_begin_nzs_of_level_after_merge_path.cc appends
            continue;
        }
        const uint64_t pos = p * work;
        uint32_t lo = 0, hi = n_rows; // first r with excl[r + 1] >= pos + 2
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if ((uint64_t)excl[mid + 1] < pos + 2) lo = mid + 1;
            else hi = mid;
        }
        part_row[p] = lo, part_nz[p] = (uint32_t)(pos - ordinal[lo]);
    }
}
} // namespace

// [sort_operator +] merge_path_{thread,warp,tblock}_operator(work): GLOBAL nz_* (sorted when asked), `level` first_row_indices_without_ending and
// first_nz_indices (with the closing nnz)
cudaError_t transform_merge_path_chain(const coo_dev &m, bool sort, uint32_t work, const char *level, bool half, double pad_bound, cudaStream_t s, transform_out &out)
{
    if (work < 1) {
        out.error = "merge path: work_size must be >= 1";
        return cudaErrorInvalidValue;
    }
    TRY(transform_row_chain(m, sort, 0, half, "ROW_TMP", pad_bound, s, out)); // the (sorted) stream and its row pointer; no padding
    uint32_t *rowptr = nullptr;
    for (size_t i = 0; i < out.arrays.size();) {
        if (out.arrays[i].pos == "ROW_TMP") {
            if (out.arrays[i].name == "first_nz_indices") rowptr = (uint32_t *)out.arrays[i].ptr;
            else cudaFreeAsync(out.arrays[i].ptr, s);
            out.arrays.erase(out.arrays.begin() + i);
        } else i++;
    }
    if (!rowptr) return cudaErrorInvalidValue;
    scratch sc(s);
    sc.bufs.push_back(rowptr); // scaffolding: freed with the scratch
    const uint32_t n_rows = (uint32_t)out.rows_kept, nnz = (uint32_t)out.stored_nnz;
    uint32_t *term = nullptr, *ordinal = nullptr, total = 0;
    TRY(sc.get(&term, (uint64_t)n_rows + 1));
    TRY(sc.get(&ordinal, (uint64_t)n_rows + 1));
    k_mp_path_terms<<<blocks_for(n_rows, 256 * 4), 256, 0, s>>>(rowptr, n_rows, term, ordinal);
    TRY(cudaGetLastError());
    TRY(scan_exclusive(term, term, n_rows, sc, s));
    TRY(scan_exclusive(ordinal, ordinal, n_rows, sc, s));
    TRY(cudaMemcpyAsync(&total, term + n_rows, 4, cudaMemcpyDeviceToHost, s));
    TRY(cudaStreamSynchronize(s));
    if (total < 2) { // no non-zero at all
        out.error = "merge path over a matrix without non-zeros";
        return cudaErrorInvalidValue;
    }
    const uint64_t count = (uint64_t)total - 1; // the first non-empty row carries no +1
    const uint32_t n_parts = (uint32_t)((count + work - 1) / work);
    uint32_t *part_row = nullptr, *part_nz = nullptr;
    TRY(alloc_out(&part_row, n_parts, s));
    take(out, level, "first_row_indices_without_ending", part_row, n_parts, GS_U32, n_rows ? n_rows - 1 : 0, 0);
    TRY(alloc_out(&part_nz, (uint64_t)n_parts + 1, s));
    take(out, level, "first_nz_indices", part_nz, (uint64_t)n_parts + 1, GS_U32, nnz, 0);
    k_mp_parts<<<blocks_for((uint64_t)n_parts + 1, 256), 256, 0, s>>>(term, ordinal, n_rows, work, n_parts, nnz, part_row, part_nz);
    TRY(cudaGetLastError());
    return cudaStreamSynchronize(s);
}

} // namespace gsd
