// B200 nnz kernel (GS_FAMILY_B200_NNZ): the nnz-direction chains (thread_bit_map / warp_segment:
// fixed 32 non-zeros per BMT, row-start bit maps) for wide dense operands, N a multiple of 32.
//
// What it replaces: thread_bit_map_reduce_to_two_register_token.cc:819-1616 gives every thread one
// BMT and one dense column group, so the col/val loads of a warp are strided by the BMT size and every
// row end is an atomicAdd.  Here the warp is the unit: it owns a RUN of consecutive BMTs, lane l owns
// dense columns l, l+32, ... (a row of B is one or more whole 128-byte lines per warp request), and
//   * a BMT's 32 column ids / values are one coalesced 128-byte load each, handed round by shuffles;
//   * the row-start bit map is warp-uniform, so the flush branch never diverges;
//   * a row that starts and ends inside the run is stored with plain coalesced stores; only the first
//     and the last partial row of a run go through red.global.add (C is zeroed per launch like for the
//     reference's bit-map families, which also covers empty rows);
//   * sixteen gathers of B (N=32; eight per column slice above) are in flight per warp before the first
//     FMA of a group; the next BMT's col/val lines are requested before the current BMT is multiplied.
// Bound: the gathers of B (stored nnz x N x sizeof(value) bytes through L2, hot columns served by
// L1/L2), not the A stream; see DESIGN.md section 4.4.
#include "common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <vector>

#include "../../../include/gsparse_b200.h"

namespace gsd {

namespace {

constexpr int RUN = 16;     // BMTs per warp (default; at most 31: one metadata lane per BMT plus the end offset)
constexpr int WARPS = 8;    // warps per CTA

__device__ __forceinline__ uint32_t out_row(const dev_arrays &a, uint32_t row) { return a.perm ? a.perm[row] + a.begin_row : row + a.begin_row; }

__device__ __forceinline__ void stg_policy(float *a, float v, uint64_t pol);
__device__ __forceinline__ void stg_policy(__half *, float, uint64_t) {}

template <typename ValT, int NV>
__device__ __forceinline__ void flush_row(ValT *C, int N, uint32_t row, int lane, float (&acc)[NV], bool atomic, uint64_t store_pol = 0)
{
    ValT *dst = C + (size_t)row * N + lane;
#pragma unroll
    for (int v = 0; v < NV; v++) {
        if (atomic) {
            if (acc[v] != 0.0f) atomic_add_out(dst, (size_t)(32 * v), acc[v]);
        } else if (store_pol != 0 && sizeof(ValT) == 4) {
            stg_policy(dst + 32 * v, acc[v], store_pol); // C is written once and never read: it must not displace B's popular rows
        } else {
            dst[32 * v] = from_f32<ValT>(acc[v]);
        }
        acc[v] = 0.0f;
    }
}

// L2 residency control for matrices whose dense operand does not fit L2 (C4: B = 1.28 GB against 126 MB).  ncu on the plain kernel
// (profiles/r2_c4_nnz_summary.md): 23.8 GB of DRAM traffic per launch for 4.28 GB of compulsory bytes, L2 read hit rate 12 % -- the
// 24.6 GB of row gathers miss because the A stream, the C stores and the cold rows of B wash the popular rows out.  With HOT the plan
// marks the columns that carry most non-zeros (bit 31 of the stored id, as many as fit GS_NNZ_HOT_MB of B) : their rows are loaded
// with an evict_last policy, every other row and the streams with evict_first, so the popular rows stay.
// MEASURED on C4, one B200 (profiles/r2_c4_nnz_summary.md): plain kernel 5.36 ms; GS_NNZ_HOT=1 (evict_first on the A, metadata and C streams
// only) 6.26 ms; GS_NNZ_HOT=2 (that + evict_last / evict_first per gathered row, 31 % of the gathers marked hot) 7.58-7.86 ms whatever the
// size of the hot set.  The hardware's own replacement does better than either hint, so both stay opt-in experiments.
__device__ __forceinline__ uint64_t policy_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float ldg_policy(const float *a, uint64_t pol)
{
    float x;
    asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(x) : "l"(a), "l"(pol));
    return x;
}
__device__ __forceinline__ float ldg_policy_na(const float *a, uint64_t pol)
{
    float x;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(x) : "l"(a), "l"(pol));
    return x;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t *a, uint64_t pol)
{
    uint32_t x;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(x) : "l"(a), "l"(pol));
    return x;
}
__device__ __forceinline__ void stg_policy(float *a, float v, uint64_t pol) { asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(a), "f"(v), "l"(pol) : "memory"); }

template <typename ColT, typename ValT, int NV, int MINB, int GROUP, int HOT = 0> __global__ void __launch_bounds__(WARPS * 32, MINB) k_b200_nnz(launch_params p, const uint32_t run)
{
    static_assert(!HOT || (sizeof(ColT) == 4 && sizeof(ValT) == 4), "the hot-column variant is built for u32 ids and fp32 values");
    uint64_t pol_last = 0, pol_first = 0;
    if (HOT) pol_last = policy_evict_last(), pol_first = policy_evict_first();
    const dev_arrays &a = p.a;
    const int lane = threadIdx.x & 31;
    const uint32_t w = blockIdx.x * WARPS + (threadIdx.x >> 5);
    const uint32_t b0 = w * run;
    if (b0 >= a.n_t) return;
    const uint32_t nb = min(run, a.n_t - b0);
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val;
    const ValT *B = (const ValT *)p.B;
    ValT *C = (ValT *)p.C;
    const int N = p.N;

    // the run's metadata: one BMT per lane, handed round by shuffles afterwards
    uint32_t m_fn = 0, m_row = 0, m_lo = 0, m_hi = 0, m_gap = 0;
    if ((uint32_t)lane <= nb) m_fn = p.first_nz_step ? (b0 + lane) * p.first_nz_step : a.t_first_nz[b0 + lane];
    if ((uint32_t)lane < nb) {
        uint64_t map = a.t_bit_map[b0 + lane];
        m_lo = (uint32_t)map, m_hi = (uint32_t)(map >> 32);
        m_row = a.t_first_row[b0 + lane];
        uint32_t b = b0 + lane;
        m_gap = (uint32_t)((a.t_seg_empty_flag[b / p.vector_width] >> (b % p.vector_width)) & 1ull);
    }

    float acc[NV];
#pragma unroll
    for (int v = 0; v < NV; v++) acc[v] = 0.0f;
    uint32_t row = __shfl_sync(0xffffffffu, m_row, 0);
    bool edge = true; // the open segment may have begun before this run: it goes to C through atomics

    uint32_t lo = __shfl_sync(0xffffffffu, m_fn, 0), hi = __shfl_sync(0xffffffffu, m_fn, 1);
    uint32_t c_cur = 0;
    float v_cur = 0.0f;
    if (lo + lane < hi) {
        if (HOT) c_cur = ldg_stream_u32((const uint32_t *)col + lo + lane, pol_first), v_cur = __uint_as_float(ldg_stream_u32((const uint32_t *)val + lo + lane, pol_first));
        else c_cur = col[lo + lane], v_cur = to_f32(val[lo + lane]);
    }

    for (uint32_t i = 0; i < nb; i++) {
        // request the next BMT's first 32 ids / values before multiplying this one
        uint32_t lo_n = hi, hi_n = __shfl_sync(0xffffffffu, m_fn, (i + 2) & 31);
        uint32_t c_nxt = 0;
        float v_nxt = 0.0f;
        if (i + 1 < nb && lo_n + lane < hi_n) {
            if (HOT) c_nxt = ldg_stream_u32((const uint32_t *)col + lo_n + lane, pol_first), v_nxt = __uint_as_float(ldg_stream_u32((const uint32_t *)val + lo_n + lane, pol_first));
            else c_nxt = col[lo_n + lane], v_nxt = to_f32(val[lo_n + lane]);
        }

        const uint32_t map_lo = __shfl_sync(0xffffffffu, m_lo, i), map_hi = __shfl_sync(0xffffffffu, m_hi, i);
        const uint32_t frow = __shfl_sync(0xffffffffu, m_row, i);
        const bool gaps = __shfl_sync(0xffffffffu, m_gap, i) != 0;
        if (frow != row) { // the previous BMT ended its row exactly at the boundary
            flush_row<ValT, NV>(C, N, out_row(a, row), lane, acc, edge, pol_first);
            edge = false;
            row = frow;
        }
        uint32_t seg = 0;
        const uint32_t cnt = hi - lo; // <= 64 (u64 bit map)
        if (cnt == 32) { // what fixed_interval_nnz_direction_thread_blocking(32) + nnz padding produce: fully unrolled
#pragma unroll
            for (int k0 = 0; k0 < 32; k0 += GROUP) {
                float x[GROUP][NV];
#pragma unroll
                for (int j = 0; j < GROUP; j++) {
                    uint32_t c = __shfl_sync(0xffffffffu, c_cur, k0 + j);
                    if (HOT == 2) { // one load instruction, the policy picked per id (no branch: the gathers of a group stay in flight together)
                        const uint64_t pol = (c >> 31) ? pol_last : pol_first;
                        const float *src = (const float *)B + (size_t)(c & 0x7fffffffu) * N + lane;
#pragma unroll
                        for (int v = 0; v < NV; v++) x[j][v] = ldg_policy(src + 32 * v, pol);
                        continue;
                    }
                    const ValT *src = B + (size_t)c * N + lane;
#pragma unroll
                    for (int v = 0; v < NV; v++) x[j][v] = to_f32(__ldg(src + 32 * v));
                }
#pragma unroll
                for (int j = 0; j < GROUP; j++) {
                    const int k = k0 + j;
                    float vv = __shfl_sync(0xffffffffu, v_cur, k);
                    if (k > 0 && ((map_lo >> k) & 1u)) {
                        flush_row<ValT, NV>(C, N, out_row(a, row), lane, acc, edge, pol_first);
                        edge = false;
                        seg++;
                        row = gaps ? frow + a.t_seg_empty_rows[a.t_seg_ptr[b0 + i] + seg] : frow + seg;
                    }
#pragma unroll
                    for (int v = 0; v < NV; v++) acc[v] = fmaf(vv, x[j][v], acc[v]);
                }
            }
        } else { // ragged BMTs (the last one of a matrix, or chains with another interval): one non-zero at a time
            const uint64_t map = ((uint64_t)map_hi << 32) | map_lo;
            for (uint32_t k = 0; k < cnt; k++) {
                if (k == 32) { // BMTs wider than 32 non-zeros: the second half is loaded in place
                    c_cur = 0, v_cur = 0.0f;
                    if (lo + 32 + lane < hi) c_cur = (uint32_t)col[lo + 32 + lane], v_cur = to_f32(val[lo + 32 + lane]);
                }
                uint32_t c = __shfl_sync(0xffffffffu, c_cur, k & 31);
                if (HOT == 2) c &= 0x7fffffffu;
                float vv = __shfl_sync(0xffffffffu, v_cur, k & 31);
                const ValT *src = B + (size_t)c * N + lane;
                if (k > 0 && ((map >> k) & 1ull)) {
                    flush_row<ValT, NV>(C, N, out_row(a, row), lane, acc, edge, pol_first);
                    edge = false;
                    seg++;
                    row = gaps ? frow + a.t_seg_empty_rows[a.t_seg_ptr[b0 + i] + seg] : frow + seg;
                }
#pragma unroll
                for (int v = 0; v < NV; v++) acc[v] = fmaf(vv, to_f32(__ldg(src + 32 * v)), acc[v]);
            }
        }
        c_cur = c_nxt, v_cur = v_nxt, lo = lo_n, hi = hi_n;
    }
    // the last open segment may continue in the next run
    flush_row<ValT, NV>(C, N, out_row(a, row), lane, acc, true);
}

unsigned nnz_run()
{
    static const int r = getenv("GS_NNZ_RUN") ? atoi(getenv("GS_NNZ_RUN")) : RUN;
    return (unsigned)(r < 1 ? 1 : r > 31 ? 31 : r);
}

template <typename ColT, typename ValT> cudaError_t launch_nv(const launch_params &p, cudaStream_t s, int hot = 0)
{
    static const int occ = getenv("GS_NNZ_OCC") ? atoi(getenv("GS_NNZ_OCC")) : 0;     // timing experiments (N=32 only)
    static const int grp = getenv("GS_NNZ_GROUP") ? atoi(getenv("GS_NNZ_GROUP")) : 0;
    const unsigned run = nnz_run();
    unsigned warps = (p.a.n_t + run - 1) / run;
    unsigned grid = (warps + WARPS - 1) / WARPS;
    if (grid == 0) return cudaSuccess;
    if (hot) { // u32 ids with the hot bit, fp32 values (b200_nnz_prepare)
        if (sizeof(ColT) != 4 || sizeof(ValT) != 4) return cudaErrorInvalidValue;
        switch (p.N / 32) {
        case 1:
            if (hot == 2) k_b200_nnz<uint32_t, float, 1, 5, 16, 2><<<grid, WARPS * 32, 0, s>>>(p, run);
            else k_b200_nnz<uint32_t, float, 1, 5, 16, 1><<<grid, WARPS * 32, 0, s>>>(p, run);
            break;
        case 2:
            if (hot == 2) k_b200_nnz<uint32_t, float, 2, 3, 8, 2><<<grid, WARPS * 32, 0, s>>>(p, run);
            else k_b200_nnz<uint32_t, float, 2, 3, 8, 1><<<grid, WARPS * 32, 0, s>>>(p, run);
            break;
        case 4:
            if (hot == 2) k_b200_nnz<uint32_t, float, 4, 2, 8, 2><<<grid, WARPS * 32, 0, s>>>(p, run);
            else k_b200_nnz<uint32_t, float, 4, 2, 8, 1><<<grid, WARPS * 32, 0, s>>>(p, run);
            break;
        default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    switch (p.N / 32) {
    case 1:
        if (grp == 8) k_b200_nnz<ColT, ValT, 1, 5, 8><<<grid, WARPS * 32, 0, s>>>(p, run);
        else if (occ == 4) k_b200_nnz<ColT, ValT, 1, 4, 16><<<grid, WARPS * 32, 0, s>>>(p, run);
        else k_b200_nnz<ColT, ValT, 1, 5, 16><<<grid, WARPS * 32, 0, s>>>(p, run);
        break;
    case 2: k_b200_nnz<ColT, ValT, 2, 3, 8><<<grid, WARPS * 32, 0, s>>>(p, run); break;
    case 4: k_b200_nnz<ColT, ValT, 4, 2, 8><<<grid, WARPS * 32, 0, s>>>(p, run); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

} // namespace

bool b200_nnz_supported(const launch_params &p)
{
    int nv = p.N / 32;
    return p.N % 32 == 0 && (nv == 1 || nv == 2 || nv == 4);
}

// Plan-time state of the hot-column variant: a copy of the column ids with bit 31 set on the columns that carry the most non-zeros, as many
// as fit `GS_NNZ_HOT_MB` (default 72) megabytes of B.  Only built when B is larger than L2 can hold (else every row stays anyway), for
// u32 ids and fp32 values; GS_NNZ_HOT=0 switches it off.  Returns cudaSuccess with *out == nullptr when the variant does not apply.
struct b200_nnz_plan {
    int mode = 0; // 1: the streams (A, metadata, C) carry an evict_first policy; 2: and B's rows evict_last on the hot columns, evict_first elsewhere
    uint32_t *col_hot = nullptr;
    uint64_t hot_cols = 0, hot_nnz = 0;
};

cudaError_t b200_nnz_prepare(const launch_params &p, b200_nnz_plan **out, cudaStream_t s)
{
    *out = nullptr;
    static const int on = getenv("GS_NNZ_HOT") ? atoi(getenv("GS_NNZ_HOT")) : 0; // off by default: both modes measured SLOWER on C4 (see below)
    static const int mb = getenv("GS_NNZ_HOT_MB") ? atoi(getenv("GS_NNZ_HOT_MB")) : 72;
    const uint64_t b_bytes = (uint64_t)p.a.cols * p.N * 4;
    if (!on || p.a.col_is_u16 || p.a.val_is_f16 || b_bytes <= 96ull * 1000 * 1000 || p.a.cols >= 0x80000000u || p.a.nnz == 0) return cudaSuccess;
    if (on == 1) { // streams only: no copy of the ids
        b200_nnz_plan *pl1 = new (std::nothrow) b200_nnz_plan;
        if (!pl1) return cudaErrorMemoryAllocation;
        pl1->mode = 1;
        *out = pl1;
        return cudaSuccess;
    }
    std::vector<uint32_t> col(p.a.nnz), cnt(p.a.cols, 0);
    cudaError_t e = cudaMemcpyAsync(col.data(), p.a.col, p.a.nnz * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return e;
    for (uint32_t c : col) {
        if (c >= p.a.cols) return cudaErrorInvalidValue;
        cnt[c]++;
    }
    const uint64_t max_hot = (uint64_t)mb * 1000 * 1000 / ((uint64_t)p.N * 4);
    if (max_hot == 0 || max_hot >= p.a.cols) return cudaSuccess;
    std::vector<uint32_t> sorted(cnt);
    std::nth_element(sorted.begin(), sorted.begin() + (sorted.size() - max_hot), sorted.end());
    const uint32_t thr = std::max<uint32_t>(2u, sorted[sorted.size() - max_hot]); // a column used once gains nothing from staying
    b200_nnz_plan *pl = new (std::nothrow) b200_nnz_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    pl->mode = 2;
    uint64_t budget = max_hot;
    std::vector<uint8_t> hot(p.a.cols, 0);
    for (uint32_t c = 0; c < p.a.cols && budget; c++)
        if (cnt[c] > thr) hot[c] = 1, budget--, pl->hot_cols++;
    for (uint32_t c = 0; c < p.a.cols && budget; c++)
        if (cnt[c] == thr) hot[c] = 1, budget--, pl->hot_cols++;
    for (uint32_t &c : col)
        if (hot[c]) c |= 0x80000000u, pl->hot_nnz++;
    if ((e = cudaMalloc((void **)&pl->col_hot, p.a.nnz * 4 + 16)) != cudaSuccess || (e = cudaMemcpyAsync(pl->col_hot, col.data(), p.a.nnz * 4, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
        (e = cudaStreamSynchronize(s)) != cudaSuccess) {
        b200_nnz_free(pl);
        return e;
    }
    *out = pl;
    return cudaSuccess;
}

void b200_nnz_free(b200_nnz_plan *pl)
{
    if (!pl) return;
    cudaFree(pl->col_hot);
    delete pl;
}

cudaError_t b200_nnz_launch(const launch_params &p0, const b200_nnz_plan *pl, cudaStream_t s)
{
    launch_params p = p0;
    if (!b200_nnz_supported(p)) return cudaErrorInvalidValue;
    if (pl && pl->mode == 2 && pl->col_hot) {
        p.a.col = pl->col_hot;
        return launch_nv<uint32_t, float>(p, s, 2);
    }
    if (pl && pl->mode == 1 && !p.a.col_is_u16 && !p.a.val_is_f16) return launch_nv<uint32_t, float>(p, s, 1);
    if (p.a.col_is_u16) return p.a.val_is_f16 ? launch_nv<uint16_t, __half>(p, s) : launch_nv<uint16_t, float>(p, s);
    return p.a.val_is_f16 ? launch_nv<uint32_t, __half>(p, s) : launch_nv<uint32_t, float>(p, s);
}

const char *b200_nnz_kernel_name(const launch_params &p, const b200_nnz_plan *pl)
{
    static thread_local char name[160];
    char hot[80] = "";
    if (pl && pl->mode == 1) snprintf(hot, sizeof hot, ",L2 evict_first on the A, metadata and C streams");
    if (pl && pl->mode == 2 && pl->col_hot) snprintf(hot, sizeof hot, ",L2 evict_last on %llu hot columns (%.0f%% of the gathers)", (unsigned long long)pl->hot_cols, 100.0 * (double)pl->hot_nnz / (double)p.a.nnz);
    snprintf(name, sizeof name, "k_b200_nnz<N%d,%s,%s,run %d BMTs/warp,%d gathers in flight%s>", p.N, p.a.col_is_u16 ? "u16" : "u32",
             p.a.val_is_f16 ? "f16" : "f32", (int)nnz_run(), p.N == 32 ? 16 : 8, hot);
    return name;
}

} // namespace gsd
