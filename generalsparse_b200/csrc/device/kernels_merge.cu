// B200 range kernel (GS_FAMILY_B200_MERGE): the device consumer of the chains whose blocks are RANGES OF THE NON-ZERO STREAM that do
// not line up with rows -- merge-path parts (merge_path_{thread,warp,tblock}_operator: get_begin_{rows,nzs}_of_level_after_merge_path.cc),
// nnz-balanced row groups (balanced_interval_row_direction_*: data_transform_common.cc:934-988), column-direction row chunks
// (fixed_interval_col_direction_*_blocking_operator) and row blocks of more than one row.  The reference has no kernel token for the
// first two (its seven drivers never reach them); this is the B200 one for all four.
//
// Work unit: a SEGMENT [a, b) of the stream, walked by a GROUP of LG lanes (LG = min(N, 32)); lane j of the group owns dense columns
// j, j + 32, ... so a gathered row of B is one coalesced request of N values (whole 128-byte lines from N = 32 up, one 32-byte sector
// at N = 8 fp32).  Segments are the chain's own blocks; a block longer than a cap is cut into equal slices at plan time (the hub row of
// a power-law matrix becomes thousands of slices), the row each slice starts in found by binary search in the row pointer -- host
// side, once.  The group loads LG column ids / values per step with one coalesced request and hands them round by shuffles; row
// ends come from the row pointer (a warp-uniform scalar per group).  A row that lies wholly inside a segment is stored plainly; the
// (at most two) rows a segment shares with its neighbours go to C through red.global.add, and C is zeroed per launch, which also
// covers empty rows.  Bound: HBM (A streamed once) while B's working set sits in L1/L2.
#include "common.cuh"
#include "transform.hpp"

#include "../../../include/gsparse_b200.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace gsd {

struct b200_merge_plan {
    uint32_t *seg_nz = nullptr, *seg_row = nullptr, *rowptr = nullptr;
    uint32_t n_seg = 0, n_rows = 0;
    uint64_t part_bytes = 0; // the chain's own range array, as the kernel's format bytes count it
    int N = 0;
    uint32_t *head_s0 = nullptr, *tail_s1 = nullptr, *ticket = nullptr, *zero_rows = nullptr; // ticket mode (see the chunk kernel)
    float *partial = nullptr;
    uint32_t n_zero = 0, max_sharers = 1;
    uint32_t cap = 0;        // no segment is longer
    bool stage = false;      // chunk kernel: the CTA's piece of the stream goes through shared memory
    bool per_thread = false; // rows of B up to 256 bytes: one lane per 16-byte chunk of the row (k_b200_merge_chunk), else a group of 32 lanes
    char name[160] = {0};
};

namespace {

constexpr int MERGE_WARPS = 8, BATCH = 8;

struct merge_args {
    const void *col, *val, *B;
    void *C;
    const uint32_t *seg_nz, *seg_row, *rowptr, *perm;
    uint32_t n_seg, begin_row;
    int N;
    uint32_t nnz, stage_col_words; // staged chunk kernel: length of the stream, words of shared memory the ids take
    // ticket mode (chunk kernel): rows shared by several segments are summed by the LAST segment to arrive, from partial sums in `partial`
    const uint32_t *head_s0, *tail_s1; // first segment that shares segment s's first row / last segment that shares its last row
    uint32_t *ticket;                  // one counter per segment (indexed by the first sharer of a row), zero between launches
    float *partial;                    // [segment][first row, last row][N]
    const uint32_t *zero_rows;         // rows with no stored non-zero: written as zeros by the kernel itself
    uint32_t n_zero;
};

template <typename ValT, int NV, int LG> __device__ __forceinline__ void flush(ValT *C, int N, uint32_t out_row, int j, float (&acc)[NV], bool edge)
{
    ValT *dst = C + (size_t)out_row * N + j;
#pragma unroll
    for (int v = 0; v < NV; v++) {
        if (edge) {
            if (acc[v] != 0.0f) atomic_add_out(dst, (size_t)(LG * v), acc[v]);
        } else dst[LG * v] = from_f32<ValT>(acc[v]);
        acc[v] = 0.0f;
    }
}


// ---- ticket mode ------------------------------------------------------------------------------------------------------------------------------
// No memset of C and no atomics on it.  A segment stores the rows that lie wholly inside it; for a row it shares with its neighbours it writes
// its partial sum (fp32) to `partial`, fences, and takes a ticket from the row's counter; the segment that draws the last ticket reads every
// sharer's partial in stream order and stores the row.  The order of the sum is fixed, so C is bit-reproducible from launch to launch (red.add
// is not), a HALF row is rounded ONCE (red.add rounds every slice's partial to fp16: 2^-11 of the partial, which breaks SURVEY 8(d)'s gate
// where slices cancel), and the launch is one kernel (a memset in front costs ~4 us of launch latency: scratch/launch_floor.sh).  The counters
// are reset by the last arriver.  Rows without non-zeros come from a list.  The plan falls back to memset + red.add when a row is shared by
// more than GS_MERGE_TICKET_MAX segments (a hub row cut into thousands: the last arriver would sum them alone).
// Lane j of a group of LG holds NC columns: j + LG * t (STRIDED, the group kernel) or j * NC + t (the chunk kernel).
template <int LG, int NC, bool STRIDED, typename ValT>
__device__ __forceinline__ void ticket_flush(const merge_args &p, uint32_t g, uint32_t r, uint32_t s0, uint32_t s1, uint32_t slot, int j, unsigned gmask, const float (&acc)[NC], ValT *dst_row)
{
    constexpr int N = LG * NC;
    auto colof = [&](int t) { return STRIDED ? j + LG * t : j * NC + t; };
    float *mine = p.partial + ((size_t)g * 2 + slot) * N;
#pragma unroll
    for (int t = 0; t < NC; t++) __stcg(mine + colof(t), acc[t]);
    __threadfence();
    unsigned tk = 0;
    if (LG > 1) __syncwarp(gmask);
    if (j == 0) tk = atomicAdd(p.ticket + s0, 1u);
    if (LG > 1) tk = __shfl_sync(gmask, tk, (threadIdx.x & 31) & ~(LG - 1));
    if (tk != s1 - s0) return;
    // every other sharer has published its partial
    __threadfence();
    if (j == 0) p.ticket[s0] = 0;
    float sum[NC];
#pragma unroll
    for (int t = 0; t < NC; t++) sum[t] = 0.0f;
    const uint32_t slot0 = p.seg_row[s0] == r ? 0u : 1u; // the first sharer holds the row as its last row unless the row begins with it
#pragma unroll 4
    for (uint32_t q = s0; q <= s1; q++) {
        const float *src = p.partial + ((size_t)q * 2 + (q == s0 ? slot0 : 0u)) * N;
#pragma unroll
        for (int t = 0; t < NC; t++) sum[t] += __ldcg(src + colof(t));
    }
    if (!STRIDED && NC * (int)sizeof(ValT) == 16) {
        uint4 o;
        ValT *ov = (ValT *)&o;
#pragma unroll
        for (int t = 0; t < NC; t++) ov[t] = from_f32<ValT>(sum[t]);
        *(uint4 *)(dst_row + j * NC) = o;
    } else {
#pragma unroll
        for (int t = 0; t < NC; t++) dst_row[colof(t)] = from_f32<ValT>(sum[t]);
    }
}

// rows without stored non-zeros (ticket mode has no memset): every thread of the grid takes its share
template <typename ValT> __device__ __forceinline__ void zero_listed_rows(const merge_args &p, int N)
{
    const uint64_t total = (uint64_t)p.n_zero * N;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t zr = p.zero_rows[i / N];
        ((ValT *)p.C)[(size_t)((p.perm ? p.perm[zr] : zr) + p.begin_row) * N + i % N] = from_f32<ValT>(0.0f);
    }
}

template <typename ColT, typename ValT, int LG, int NV, bool TICKET> __global__ void __launch_bounds__(MERGE_WARPS * 32) k_b200_merge(merge_args p)
{
    const uint32_t tid = blockIdx.x * (MERGE_WARPS * 32) + threadIdx.x;
    const uint32_t g = tid / LG;
    const int j = tid % LG, lane = threadIdx.x & 31;
    if (TICKET) zero_listed_rows<ValT>(p, p.N);
    if (g >= p.n_seg) return; // whole groups leave together
    const unsigned mask = LG == 32 ? 0xffffffffu : (((1u << LG) - 1u) << (lane & ~(LG - 1)));
    const ColT *col = (const ColT *)p.col;
    const ValT *val = (const ValT *)p.val, *B = (const ValT *)p.B;
    ValT *C = (ValT *)p.C;
    const int N = p.N;
    const uint32_t a = p.seg_nz[g], b = p.seg_nz[g + 1];
    uint32_t r = p.seg_row[g];
    uint32_t row_begin = p.rowptr[r], e = p.rowptr[r + 1];
    bool edge = a != row_begin; // the first row began before this segment: others add to it too
    bool first = true;          // still in the segment's first row
    const uint32_t head_s0 = TICKET ? p.head_s0[g] : g, tail_s1 = TICKET ? p.tail_s1[g] : g;
    float acc[NV];
#pragma unroll
    for (int v = 0; v < NV; v++) acc[v] = 0.0f;
    auto out = [&](uint32_t row) { return (p.perm ? p.perm[row] : row) + p.begin_row; };
    // closing: the segment ends here (the row may go on in the next one)
    auto flush_row = [&](bool as_edge, bool closing) {
        if (TICKET && as_edge) {
            ticket_flush<LG, NV, true, ValT>(p, g, r, first ? head_s0 : g, closing && b != e ? tail_s1 : g, first ? 0u : 1u, j, mask, acc, C + (size_t)out(r) * N);
#pragma unroll
            for (int v = 0; v < NV; v++) acc[v] = 0.0f;
        } else flush<ValT, NV, LG>(C, N, out(r), j, acc, as_edge);
    };

    for (uint32_t k = a; k < b; k += LG) {
        uint32_t c_mine = 0;
        float v_mine = 0.0f;
        if (k + j < b) c_mine = (uint32_t)col[k + j], v_mine = to_f32(val[k + j]);
        const uint32_t cnt = min((uint32_t)LG, b - k);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < cnt; j0 += BATCH) {
            float x[BATCH][NV];
            // the gathers of a batch are in flight together; entries past the end repeat the lane's own id with value 0
#pragma unroll
            for (int q = 0; q < BATCH; q++) {
                const uint32_t c = __shfl_sync(mask, c_mine, (int)((j0 + q) & (LG - 1)), LG);
                const ValT *src = B + (size_t)c * N + j;
#pragma unroll
                for (int v = 0; v < NV; v++) x[q][v] = (LG >= BATCH || j0 + q < cnt) ? to_f32(__ldg(src + LG * v)) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < BATCH; q++) {
                if (LG < BATCH && q >= LG) break;
                const uint32_t kk = k + j0 + q;
                const float vv = __shfl_sync(mask, v_mine, (int)((j0 + q) & (LG - 1)), LG);
                if (j0 + q < cnt) {
                    while (kk == e) { // the row ended (possibly followed by empty rows: C keeps its zeros)
                        if (e != row_begin) flush_row(edge, false);
                        edge = false, first = false;
                        r++, row_begin = e, e = p.rowptr[r + 1];
                    }
#pragma unroll
                    for (int v = 0; v < NV; v++) acc[v] = fmaf(vv, x[q][v], acc[v]);
                }
            }
        }
    }
    if (b > a) flush_row(edge || b != e, true); // the last row may go on in the next segment
}

// ---- narrow N: one 16-byte chunk of B's row per lane ---------------------------------------------------------------------------------------
// ncu on the group kernel above at C1 (N = 8, merge_path_thread(32); profiles/r2_c1_summary.md): 14.1 M warp instructions for 1.79 M non-zeros,
// the issue slots 52 % busy, short-scoreboard (shuffle) stalls on top -- with 8 lanes per segment and 4 bytes per lane a warp instruction
// serves 4 non-zeros, and handing ids and values round by shuffle costs more than the arithmetic.  And the one-warp-per-row gather kernel
// sits at the same 26 us for another reason: every 128-bit load whose 32 lanes touch 32 different lines costs ~2 cycles of the L1 wavefront
// queue per line (B300_MICROARCH.md, "L1tex wavefront queue"), and it issues two per non-zero for a 32-byte row of B.
// Here a segment belongs to LG = (row bytes) / 16 lanes and every lane owns ONE 16-byte chunk of the row (4 fp32 or 8 fp16 columns):
// one load instruction per non-zero and group, one L1 wavefront per gathered row, ids and values read by every lane of the group itself
// (the same address: one wavefront, no shuffle), four non-zeros in flight.  A warp instruction serves 32 / LG non-zeros (16 at N = 8 fp32).
// Rows narrower than 16 bytes (N <= 2 fp32 ...) run with one lane per segment and scalar loads.  Edges as above (red.add), whole rows stored.
template <size_t B> struct raw_of;
template <> struct raw_of<2> { typedef uint16_t type; };
template <> struct raw_of<4> { typedef uint32_t type; };
__device__ __forceinline__ void red_add_chunk(float *dst, const float (&acc)[4])
{
#pragma unroll
    for (int t = 0; t < 4; t++)
        if (acc[t] != 0.0f) atomicAdd(dst + t, acc[t]);
}
__device__ __forceinline__ void red_add_chunk(__half *dst, const float (&acc)[8])
{
#pragma unroll
    for (int t = 0; t < 4; t++) // packed: one red.global.add.noftz.f16x2 per pair of columns
        if (acc[2 * t] != 0.0f || acc[2 * t + 1] != 0.0f) atomicAdd((__half2 *)dst + t, __floats2half2_rn(acc[2 * t], acc[2 * t + 1]));
}
template <int CH> __device__ __forceinline__ void red_add_chunk(float *dst, const float (&acc)[CH])
{
#pragma unroll
    for (int t = 0; t < CH; t++)
        if (acc[t] != 0.0f) atomicAdd(dst + t, acc[t]);
}
template <int CH> __device__ __forceinline__ void red_add_chunk(__half *dst, const float (&acc)[CH])
{
#pragma unroll
    for (int t = 0; t < CH; t++)
        if (acc[t] != 0.0f) atomicAdd(dst + t, __float2half_rn(acc[t]));
}

// Staging (STAGE): the segments of a CTA are consecutive, so the CTA's piece of the stream is ONE contiguous range of at most
// (128 / LG) * cap non-zeros.  ncu on the unstaged kernel (C1, N = 8: 17.3 us, long-scoreboard 21 per issue, 45 % of the warp slots, L1 data pipe
// 53 %) shows a latency chain: every lane walks its own segment, so each round is a DRAM round trip for ids / values (16 different lines per
// load instruction) followed by an L2 round trip for B, cap / U times in a row.  Staged, the CTA copies its range to shared memory with
// coalesced 128-bit loads, all in flight at once (one DRAM round trip), and the walk reads ids and values from shared memory: what is left
// per round is the gather of B alone, eight rows in flight.  Shared memory is padded by one word per 32 so that lanes whose segments start
// a regular distance apart (every slice 32 long) fall into different banks.
template <typename T> __device__ __forceinline__ T staged_at(const uint32_t *sw, uint32_t idx)
{
    constexpr uint32_t EPW = 4 / sizeof(T); // elements per word
    const uint32_t w = idx / EPW;
    return ((const T *)(sw + w + (w >> 5)))[idx % EPW];
}
// eight elements (one 16-byte-aligned piece of the stream; `first` is a multiple of 8) from global to padded shared memory
template <typename T> __device__ __forceinline__ void stage8(uint32_t *sw, const T *g, uint32_t chunk, uint32_t first, uint32_t total)
{
    constexpr int W = 2 * (int)sizeof(T); // words per eight elements: 4 or 8, never across a 32-word boundary
    uint32_t w[W];
    if (first + 8 <= total) {
#pragma unroll
        for (int q = 0; q < W / 4; q++) {
            const uint4 x = __ldg((const uint4 *)(g + first) + q);
            w[4 * q] = x.x, w[4 * q + 1] = x.y, w[4 * q + 2] = x.z, w[4 * q + 3] = x.w;
        }
    } else { // the end of the stream: nothing is read past it
        typedef typename raw_of<sizeof(T)>::type R;
        R *e = (R *)w;
#pragma unroll
        for (int t = 0; t < 8; t++) e[t] = first + t < total ? ((const R *)g)[first + t] : (R)0;
    }
    const uint32_t w0 = chunk * W, at = w0 + (w0 >> 5);
#pragma unroll
    for (int q = 0; q < W; q++) sw[at + q] = w[q];
}
__host__ __device__ constexpr uint32_t stage_words(uint32_t elems, uint32_t elem_bytes)
{
    return (elems * elem_bytes / 4) + (elems * elem_bytes / 4 >> 5) + 8; // padded, plus slack for the last (partly filled) piece
}

// programmatic dependent launch (ticket mode only: a memset node in front is not a kernel)
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename ColT, typename ValT, int NT, int LG, bool STAGE, bool TICKET> __global__ void __launch_bounds__(128) k_b200_merge_chunk(merge_args p)
{
    grid_dep_launch(); // the next launch on the stream may begin its prologue (plan tables and the stream of A only)
    constexpr int CH = NT / LG;                               // columns per lane
    constexpr bool VEC = CH * (int)sizeof(ValT) == 16;        // the lane's piece of a row is one 128-bit load
    static_assert(NT % LG == 0 && (VEC || LG == 1), "chunking");
    extern __shared__ __align__(16) uint32_t stage_mem[];
    const uint32_t tid = blockIdx.x * 128 + threadIdx.x, g = tid / LG;
    const int j = tid % LG;
    const ColT *col = (const ColT *)p.col;
    const ValT *val = (const ValT *)p.val, *B = (const ValT *)p.B + j * CH;
    ValT *C = (ValT *)p.C + j * CH;
    const bool mine = g < p.n_seg;
    constexpr bool ticketed = TICKET;
    const uint32_t a = mine ? p.seg_nz[g] : 0u, b = mine ? p.seg_nz[g + 1] : 0u;
    uint32_t r = mine ? p.seg_row[g] : 0u;
    const uint32_t head_s0 = mine && ticketed ? p.head_s0[g] : g, tail_s1 = mine && ticketed ? p.tail_s1[g] : g;
    uint32_t base = 0;
    const uint32_t *s_col = stage_mem, *s_val = stage_mem + p.stage_col_words;
    if (STAGE) {
        constexpr uint32_t SEGS = 128 / LG;
        const uint32_t g0 = blockIdx.x * SEGS, g1 = min(g0 + SEGS, p.n_seg);
        const uint32_t A1 = p.seg_nz[g1];
        base = p.seg_nz[g0] & ~7u;
#pragma unroll 2
        for (uint32_t i = threadIdx.x; base + 8 * i < A1; i += 128) {
            stage8<ColT>(stage_mem, col, i, base + 8 * i, p.nnz);
            stage8<ValT>(stage_mem + p.stage_col_words, val, i, base + 8 * i, p.nnz);
        }
    }
    // the row pointer is read one row ahead: the load for the next row end is in flight while this row is summed
    uint32_t row_begin = p.rowptr[r], e = p.rowptr[r + 1], e_next = p.rowptr[r + 2];
    if (STAGE) __syncthreads();
    grid_dep_wait(); // B, C, the tickets and the partial sums belong to the launch before this one until it has completed
    if (ticketed) zero_listed_rows<ValT>(p, NT);
    if (!mine) return;
    const unsigned gmask = LG == 1 ? 0u : (((1u << LG) - 1u) << ((threadIdx.x & 31) & ~(LG - 1)));
    bool first = true; // still in the segment's first row
    bool edge = a != row_begin;
    float acc[CH];
#pragma unroll
    for (int t = 0; t < CH; t++) acc[t] = 0.0f;
    // closing: the segment ends here (the row may go on in the next one)
    auto flush_row = [&](bool as_edge, bool closing) {
        ValT *dst = C + (size_t)((p.perm ? p.perm[r] : r) + p.begin_row) * NT;
        if (as_edge && ticketed) {
            ticket_flush<LG, CH, false, ValT>(p, g, r, first ? head_s0 : g, closing && b != e ? tail_s1 : g, first ? 0u : 1u, j, gmask, acc, (ValT *)p.C + (size_t)((p.perm ? p.perm[r] : r) + p.begin_row) * NT);
        } else if (as_edge) red_add_chunk(dst, acc);
        else if (VEC) {
            uint4 o;
            ValT *ov = (ValT *)&o;
#pragma unroll
            for (int t = 0; t < CH; t++) ov[t] = from_f32<ValT>(acc[t]);
            *(uint4 *)dst = o;
        } else {
#pragma unroll
            for (int t = 0; t < CH; t++) dst[t] = from_f32<ValT>(acc[t]);
        }
#pragma unroll
        for (int t = 0; t < CH; t++) acc[t] = 0.0f;
    };
    constexpr int U = STAGE ? 8 : 4; // non-zeros whose loads are in flight together
    for (uint32_t k = a; k < b; k += U) {
        uint32_t c[U];
        float v[U];
        uint4 x[U];
        float xs[U][VEC ? 1 : CH];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const bool in = k + u < b;
            if (STAGE) c[u] = in ? (uint32_t)staged_at<ColT>(s_col, k + u - base) : 0u, v[u] = in ? to_f32(staged_at<ValT>(s_val, k + u - base)) : 0.0f;
            else c[u] = in ? (uint32_t)col[k + u] : 0u, v[u] = in ? to_f32(val[k + u]) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const ValT *src = B + (size_t)c[u] * NT;
            if (VEC) x[u] = __ldg((const uint4 *)src);
            else {
#pragma unroll
                for (int t = 0; t < CH; t++) xs[u][VEC ? 0 : t] = to_f32(__ldg(src + t));
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (k + u < b) {
                while (k + u == e) { // the row ended (possibly followed by empty rows)
                    if (e != row_begin) flush_row(edge, false);
                    edge = false, first = false;
                    r++, row_begin = e, e = e_next, e_next = p.rowptr[r + 2];
                }
                if (VEC) {
                    const ValT *xv = (const ValT *)&x[u];
#pragma unroll
                    for (int t = 0; t < CH; t++) acc[t] = fmaf(v[u], to_f32(xv[t]), acc[t]);
                } else {
#pragma unroll
                    for (int t = 0; t < CH; t++) acc[t] = fmaf(v[u], xs[u][VEC ? 0 : t], acc[t]);
                }
            }
        }
    }
    if (b > a) flush_row(edge || b != e, true);
}

// lanes per segment of the chunk kernel: bytes of a row of B / 16, at least 1
template <typename ValT> constexpr int chunk_lanes(int N) { return N * (int)sizeof(ValT) >= 16 ? N * (int)sizeof(ValT) / 16 : 1; }

template <typename ColT, typename ValT, bool STAGE, bool TICKET> const void *pick_thread_s(int N)
{
    switch (N) {
    case 1: return (const void *)k_b200_merge_chunk<ColT, ValT, 1, 1, STAGE, TICKET>;
    case 2: return (const void *)k_b200_merge_chunk<ColT, ValT, 2, 1, STAGE, TICKET>;
    case 4: return (const void *)k_b200_merge_chunk<ColT, ValT, 4, 1, STAGE, TICKET>;
    case 8: return (const void *)k_b200_merge_chunk<ColT, ValT, 8, chunk_lanes<ValT>(8), STAGE, TICKET>;
    case 16: return (const void *)k_b200_merge_chunk<ColT, ValT, 16, chunk_lanes<ValT>(16), STAGE, TICKET>;
    case 32: return (const void *)k_b200_merge_chunk<ColT, ValT, 32, chunk_lanes<ValT>(32), STAGE, TICKET>;
    // rows of B up to 256 bytes (fp32 N = 64, fp16 N = 64 / 128): 8 or 16 lanes per segment, unstaged (the staged form is for LG <= 4)
    case 64: return STAGE ? nullptr : (const void *)k_b200_merge_chunk<ColT, ValT, 64, chunk_lanes<ValT>(64), false, TICKET>;
    case 128: return (STAGE || sizeof(ValT) != 2) ? nullptr : (const void *)k_b200_merge_chunk<ColT, ValT, 128, (sizeof(ValT) == 2 ? 16 : 32), false, TICKET>;
    default: return nullptr;
    }
}
template <typename ColT, typename ValT> const void *pick_thread(int N, bool stage, bool ticket)
{
    return stage ? (ticket ? pick_thread_s<ColT, ValT, true, true>(N) : pick_thread_s<ColT, ValT, true, false>(N))
                 : (ticket ? pick_thread_s<ColT, ValT, false, true>(N) : pick_thread_s<ColT, ValT, false, false>(N));
}

template <typename ColT, typename ValT, bool TICKET> const void *pick_t(int N)
{
    switch (N) {
    case 1: return (const void *)k_b200_merge<ColT, ValT, 1, 1, TICKET>;
    case 2: return (const void *)k_b200_merge<ColT, ValT, 2, 1, TICKET>;
    case 4: return (const void *)k_b200_merge<ColT, ValT, 4, 1, TICKET>;
    case 8: return (const void *)k_b200_merge<ColT, ValT, 8, 1, TICKET>;
    case 16: return (const void *)k_b200_merge<ColT, ValT, 16, 1, TICKET>;
    case 32: return (const void *)k_b200_merge<ColT, ValT, 32, 1, TICKET>;
    case 64: return (const void *)k_b200_merge<ColT, ValT, 32, 2, TICKET>;
    case 128: return (const void *)k_b200_merge<ColT, ValT, 32, 4, TICKET>;
    case 256: return (const void *)k_b200_merge<ColT, ValT, 32, 8, TICKET>;
    case 512: return (const void *)k_b200_merge<ColT, ValT, 32, 16, TICKET>;
    default: return nullptr;
    }
}
template <typename ColT, typename ValT> const void *pick(int N, bool ticket = false) { return ticket ? pick_t<ColT, ValT, true>(N) : pick_t<ColT, ValT, false>(N); }

inline unsigned grid_1d(uint64_t n) { return (unsigned)std::min<uint64_t>(std::max<uint64_t>(1, (n + 255) / 256), 148u * 32u); }

int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}


// ---- the plan built on the device -----------------------------------------------------------------------------------------------------------------
// What b200_merge_prepare computes on the host -- segments, their rows, who shares which row, the rows without non-zeros -- from the chain's
// arrays as they lie on the device: the host version downloads the parts and, for merge-path and balanced chains, the row id of every stored
// non-zero (13.8 ms for the C1 stand-in, 139 ms for a 20 M-nnz block, against an 18 us / 0.5 ms kernel).  Same arrays, element for element
// (tests/test_merge_kernel_gpu.py::test_plan_built_on_device_equals_host_built).
__device__ __forceinline__ uint32_t upper_bound_u32(const uint32_t *a, uint32_t n, uint32_t x) // first index with a[i] > x
{
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (a[mid] <= x) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
__global__ void k_mp_check(const uint32_t *parts, uint32_t n_parts, const uint32_t *rows, uint32_t n_rows_ids, const uint32_t *part_rows, uint32_t *bad)
{
    const uint64_t n = max((uint64_t)n_parts, (uint64_t)n_rows_ids);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (i < n_parts && parts[i + 1] < parts[i]) *bad = 1;
        if (rows && i && i < n_rows_ids && rows[i] < rows[i - 1]) *bad = 1;
        if (part_rows && i && i < n_parts && part_rows[i] < part_rows[i - 1]) *bad = 1;
    }
}
// column-direction chunks: part i lies in row part_rows[i] (ascending); row r's non-zeros begin where its first part begins
__global__ void k_mp_rowptr_from_part_rows(const uint32_t *parts, const uint32_t *part_rows, uint32_t n_parts, uint32_t n_rows, uint32_t *rowptr)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= n_rows; r += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = n_parts; // first part whose row is >= r
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (part_rows[mid] < r) lo = mid + 1;
            else hi = mid;
        }
        rowptr[r] = parts[lo];
    }
}
__global__ void k_mp_pieces(const uint32_t *parts, uint32_t n_parts, uint32_t cap, uint32_t *pieces)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_parts; i += (uint64_t)gridDim.x * blockDim.x) pieces[i] = (parts[i + 1] - parts[i] + cap - 1) / cap;
}
// segment s is slice q of the part that holds it: equal slices, the row of its first non-zero by binary search in the row pointer
__global__ void k_mp_segments(const uint32_t *parts, const uint32_t *piece_off, uint32_t n_parts, const uint32_t *rowptr, uint32_t n_rows, uint32_t n_seg, uint32_t *seg_nz, uint32_t *seg_row)
{
    for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s <= n_seg; s += (uint64_t)gridDim.x * blockDim.x) {
        if (s == n_seg) {
            seg_nz[s] = parts[n_parts];
            continue;
        }
        const uint32_t i = upper_bound_u32(piece_off, n_parts + 1, (uint32_t)s) - 1, q = (uint32_t)s - piece_off[i], pieces = piece_off[i + 1] - piece_off[i];
        const uint32_t a = parts[i], b = parts[i + 1], lo = a + (uint32_t)((uint64_t)(b - a) * q / pieces);
        seg_nz[s] = lo, seg_row[s] = upper_bound_u32(rowptr, n_rows + 1, lo) - 1;
    }
}
__global__ void k_mp_sharers(const uint32_t *seg_nz, const uint32_t *seg_row, uint32_t n_seg, const uint32_t *rowptr, uint32_t n_rows, uint32_t *head, uint32_t *tail, uint32_t *most)
{
    uint32_t mx = 1;
    for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_seg; q += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t a = seg_nz[q], b = seg_nz[q + 1], f = seg_row[q], l = upper_bound_u32(rowptr, n_rows + 1, b - 1) - 1;
        auto seg_of = [&](uint32_t nz) { // the segment that holds stream position nz
            const uint32_t u = upper_bound_u32(seg_nz, n_seg + 1, nz);
            return min(u ? u - 1 : 0u, n_seg - 1);
        };
        const uint32_t h = rowptr[f] < a ? seg_of(rowptr[f]) : (uint32_t)q, t = rowptr[l + 1] > b ? seg_of(rowptr[l + 1] - 1) : (uint32_t)q;
        head[q] = h, tail[q] = t;
        mx = max(mx, max((uint32_t)q - h, t - (uint32_t)q) + 1);
    }
    atomicMax(most, mx);
}
__global__ void k_mp_zero_flags(const uint32_t *rowptr, uint32_t n_rows, uint32_t rows_total, uint32_t parts_end, uint32_t *flag)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows_total; r += (uint64_t)gridDim.x * blockDim.x)
        flag[r] = r >= n_rows || rowptr[r + 1] == rowptr[r] || rowptr[r] >= parts_end;
}
__global__ void k_mp_zero_list(const uint32_t *flag, const uint32_t *off, uint32_t rows_total, uint32_t *list)
{
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows_total; r += (uint64_t)gridDim.x * blockDim.x)
        if (flag[r]) list[off[r]] = (uint32_t)r;
}

struct merge_choice {
    bool per_thread;
    int LG;
    uint32_t cap;
};
merge_choice choose(const launch_params &p)
{
    merge_choice c;
    c.per_thread = (uint64_t)p.N * (p.a.val_is_f16 ? 2 : 4) <= (uint64_t)env_int("GS_MERGE_CHUNK_ROW_BYTES", 256) && env_int("GS_MERGE_THREAD", 1) != 0;
    c.LG = c.per_thread ? (p.a.val_is_f16 ? chunk_lanes<__half>(p.N) : chunk_lanes<float>(p.N)) : p.N < 32 ? p.N : 32;
    c.cap = (uint32_t)std::max(c.LG, env_int("GS_MERGE_CAP", c.per_thread ? (c.LG >= 4 ? 64 : c.LG == 2 ? 40 : 32) : c.LG == 32 ? 512 : 32 * c.LG)); // (C1 stand-in, cap 32 / 40 / 48 at two lanes: 18.3 / 16.0 / 17.7 us) // chunk kernel: every lane of a group walks the whole segment
    return c;
}

#define MP_TRY(x)                                                                                                                                    \
    do {                                                                                                                                             \
        cudaError_t e_ = (x);                                                                                                                        \
        if (e_ != cudaSuccess) return e_;                                                                                                            \
    } while (0)

// fills pl (the caller frees it on failure); cudaErrorNotSupported: let the host path decide (bad input is refused there with the same code as before)
cudaError_t prepare_on_device(const launch_params &p, const uint32_t *parts, uint32_t n_parts, const uint32_t *nz_rows, const uint32_t *part_rows, bool parts_are_rows, b200_merge_plan *pl,
                              cudaStream_t s)
{
    const uint32_t nnz = (uint32_t)p.a.nnz;
    const merge_choice ch = choose(p);
    struct tmp_bufs {
        std::vector<void *> v;
        ~tmp_bufs()
        {
            for (void *q : v) cudaFree(q);
        }
        cudaError_t get(uint32_t **q, uint64_t n)
        {
            cudaError_t e = cudaMalloc((void **)q, std::max<uint64_t>(n, 1) * 4 + 16);
            if (e == cudaSuccess) v.push_back(*q);
            return e;
        }
    } tmp;
    uint32_t *scal = nullptr, h[4] = {0, 0, 0, 0}; // [0] bad input, [1] most sharers
    MP_TRY(tmp.get(&scal, 4));
    MP_TRY(cudaMemsetAsync(scal, 0, 16, s));
    k_mp_check<<<grid_1d(std::max<uint64_t>(n_parts, nz_rows ? nnz : 0)), 256, 0, s>>>(parts, n_parts, nz_rows, nz_rows ? nnz : 0u, part_rows, scal);
    uint32_t ends[2] = {0, 0}, last_row = 0;
    MP_TRY(cudaMemcpyAsync(&ends[0], parts, 4, cudaMemcpyDeviceToHost, s));
    MP_TRY(cudaMemcpyAsync(&ends[1], parts + n_parts, 4, cudaMemcpyDeviceToHost, s));
    if (nz_rows) MP_TRY(cudaMemcpyAsync(&last_row, nz_rows + nnz - 1, 4, cudaMemcpyDeviceToHost, s));
    else if (part_rows) MP_TRY(cudaMemcpyAsync(&last_row, part_rows + n_parts - 1, 4, cudaMemcpyDeviceToHost, s));
    MP_TRY(cudaMemcpyAsync(h, scal, 4, cudaMemcpyDeviceToHost, s));
    MP_TRY(cudaStreamSynchronize(s));
    if (h[0] || ends[0] != 0 || ends[1] > nnz) return cudaErrorNotSupported;
    uint32_t n_rows = 0;
    if (nz_rows || part_rows) {
        if (last_row >= 0xfffffff0u) return cudaErrorNotSupported;
        n_rows = last_row + 1;
    } else if (parts_are_rows) n_rows = n_parts;
    else return cudaErrorNotSupported;
    MP_TRY(cudaMalloc((void **)&pl->rowptr, ((size_t)n_rows + 1) * 4 + 16));
    if (nz_rows) MP_TRY(rowptr_from_sorted_rows(nz_rows, nnz, n_rows, pl->rowptr, s));
    else if (part_rows) {
        k_mp_rowptr_from_part_rows<<<grid_1d((uint64_t)n_rows + 1), 256, 0, s>>>(parts, part_rows, n_parts, n_rows, pl->rowptr);
        MP_TRY(cudaGetLastError());
    } else MP_TRY(cudaMemcpyAsync(pl->rowptr, parts, ((size_t)n_rows + 1) * 4, cudaMemcpyDeviceToDevice, s));
    uint32_t row_end = 0;
    MP_TRY(cudaMemcpyAsync(&row_end, pl->rowptr + n_rows, 4, cudaMemcpyDeviceToHost, s));

    // segments
    pl->per_thread = ch.per_thread, pl->cap = ch.cap;
    if (ch.per_thread) {
        const uint32_t elems = (128 / ch.LG) * ch.cap + 8;
        const uint64_t bytes = 4ull * (stage_words(elems, p.a.col_is_u16 ? 2 : 4) + stage_words(elems, p.a.val_is_f16 ? 2 : 4));
        pl->stage = env_int("GS_MERGE_STAGE", ch.LG <= 4) != 0 && ch.LG <= 8 && bytes <= 48 * 1024 && (((uintptr_t)p.a.col | (uintptr_t)p.a.val) & 15) == 0;
    }
    uint32_t *piece_off = nullptr, n_seg = 0;
    MP_TRY(tmp.get(&piece_off, (uint64_t)n_parts + 1));
    k_mp_pieces<<<grid_1d(n_parts), 256, 0, s>>>(parts, n_parts, ch.cap, piece_off);
    MP_TRY(cudaGetLastError());
    MP_TRY(scan_exclusive_u32(piece_off, piece_off, n_parts, s));
    MP_TRY(cudaMemcpyAsync(&n_seg, piece_off + n_parts, 4, cudaMemcpyDeviceToHost, s));
    MP_TRY(cudaStreamSynchronize(s));
    if (row_end < ends[1] || n_seg == 0) return cudaErrorNotSupported;
    pl->n_seg = n_seg, pl->n_rows = n_rows, pl->N = p.N, pl->part_bytes = ((uint64_t)n_parts + 1) * 4;
    MP_TRY(cudaMalloc((void **)&pl->seg_nz, ((size_t)n_seg + 1) * 4 + 16));
    MP_TRY(cudaMalloc((void **)&pl->seg_row, (size_t)n_seg * 4 + 16));
    k_mp_segments<<<grid_1d((uint64_t)n_seg + 1), 256, 0, s>>>(parts, piece_off, n_parts, pl->rowptr, n_rows, n_seg, pl->seg_nz, pl->seg_row);
    MP_TRY(cudaGetLastError());

    // tickets: who shares which row (the rule of b200_merge_prepare)
    const bool ticket_default = p.a.val_is_f16 || (ch.per_thread && (uint64_t)p.N * 4 <= 128);
    if (p.a.begin_row == 0 && env_int("GS_MERGE_TICKET", ticket_default ? 1 : 0) != 0) {
        uint32_t *head = nullptr, *tail = nullptr;
        MP_TRY(cudaMalloc((void **)&head, (size_t)n_seg * 4 + 16));
        pl->head_s0 = head;
        MP_TRY(cudaMalloc((void **)&tail, (size_t)n_seg * 4 + 16));
        pl->tail_s1 = tail;
        k_mp_sharers<<<grid_1d(n_seg), 256, 0, s>>>(pl->seg_nz, pl->seg_row, n_seg, pl->rowptr, n_rows, head, tail, scal + 1);
        MP_TRY(cudaGetLastError());
        MP_TRY(cudaMemcpyAsync(h, scal, 8, cudaMemcpyDeviceToHost, s));
        MP_TRY(cudaStreamSynchronize(s));
        pl->max_sharers = h[1] ? h[1] : 1;
        if (pl->max_sharers <= (uint32_t)env_int("GS_MERGE_TICKET_MAX", 64) && (uint64_t)n_seg * 2 * p.N * sizeof(float) <= (uint64_t)env_int("GS_MERGE_TICKET_MB", 512) << 20) {
            const uint32_t rows_total = std::max(n_rows, p.a.rows);
            uint32_t *flag = nullptr, *off = nullptr, n_zero = 0;
            MP_TRY(tmp.get(&flag, rows_total));
            MP_TRY(tmp.get(&off, (uint64_t)rows_total + 1));
            k_mp_zero_flags<<<grid_1d(rows_total), 256, 0, s>>>(pl->rowptr, n_rows, rows_total, ends[1], flag);
            MP_TRY(cudaGetLastError());
            MP_TRY(scan_exclusive_u32(flag, off, rows_total, s));
            MP_TRY(cudaMemcpyAsync(&n_zero, off + rows_total, 4, cudaMemcpyDeviceToHost, s));
            MP_TRY(cudaStreamSynchronize(s));
            if (n_zero) {
                MP_TRY(cudaMalloc((void **)&pl->zero_rows, (size_t)n_zero * 4 + 16));
                k_mp_zero_list<<<grid_1d(rows_total), 256, 0, s>>>(flag, off, rows_total, pl->zero_rows);
                MP_TRY(cudaGetLastError());
            }
            pl->n_zero = n_zero;
            MP_TRY(cudaMalloc((void **)&pl->ticket, (size_t)n_seg * 4));
            MP_TRY(cudaMemsetAsync(pl->ticket, 0, (size_t)n_seg * 4, s));
            MP_TRY(cudaMalloc((void **)&pl->partial, (size_t)n_seg * 2 * p.N * sizeof(float)));
        } else {
            cudaFree(pl->head_s0), cudaFree(pl->tail_s1);
            pl->head_s0 = pl->tail_s1 = nullptr;
        }
    }
    MP_TRY(cudaStreamSynchronize(s));
    snprintf(pl->name, sizeof pl->name, "k_b200_merge%s<N%d,%s,%s,%d lane%s/segment,%s%s%u segments of <=%u nnz from %u parts>", ch.per_thread ? "_chunk" : "", p.N, p.a.col_is_u16 ? "u16" : "u32",
             p.a.val_is_f16 ? "f16" : "f32", ch.LG, ch.LG > 1 ? "s" : "", pl->stage ? "staged," : "", pl->ticket ? "tickets," : "", pl->n_seg, ch.cap, n_parts);
    return cudaSuccess;
}

} // namespace

bool b200_merge_supported(int N) { return pick<uint32_t, float>(N) != nullptr; }

// parts: the chain's range array on the DEVICE (n_parts + 1 offsets into the stream); rows_dev: nz_row_indices on the device (u32, one per
// stored non-zero) or null, in which case chunk_rows (row of every part, col-direction chains) or the parts themselves (one row per part)
// give the row pointer.
cudaError_t b200_merge_prepare(const launch_params &p, const uint32_t *parts_dev, uint32_t n_parts, const uint32_t *nz_rows_dev, const uint32_t *part_rows_dev, bool parts_are_rows,
                               b200_merge_plan **out, cudaStream_t s)
{
    *out = nullptr;
    if (!b200_merge_supported(p.N) || n_parts == 0 || p.a.nnz == 0 || p.a.nnz > 0xfffffff0ull) return cudaErrorInvalidValue;
    b200_merge_plan *pl = new (std::nothrow) b200_merge_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    auto fail = [&](cudaError_t e) {
        b200_merge_free(pl);
        return e;
    };
    cudaError_t e;
    if (env_int("GS_MERGE_DEVICE_PREPARE", 1) != 0) { // the plan built by device passes; anything it does not take goes through the host code below
        e = prepare_on_device(p, parts_dev, n_parts, nz_rows_dev, part_rows_dev, parts_are_rows, pl, s);
        if (e == cudaSuccess) {
            *out = pl;
            return cudaSuccess;
        }
        cudaGetLastError(); // input the passes do not take, or a pass that failed (out of memory ...): the host code decides
        b200_merge_free(pl);
        if (!(pl = new (std::nothrow) b200_merge_plan)) return cudaErrorMemoryAllocation;
    }
    std::vector<uint32_t> parts(n_parts + 1), rowptr;
    if ((e = cudaMemcpyAsync(parts.data(), parts_dev, parts.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    const uint32_t nnz = (uint32_t)p.a.nnz;
    uint32_t n_rows = 0;
    if (nz_rows_dev) { // COO row ids of the (padded, possibly sorted) stream -> row pointer
        std::vector<uint32_t> rows(nnz);
        if ((e = cudaMemcpyAsync(rows.data(), nz_rows_dev, (size_t)nnz * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = rows[nnz - 1] + 1;
        rowptr.assign((size_t)n_rows + 1, 0);
        for (uint32_t i = 0; i < nnz; i++) {
            if (rows[i] >= n_rows || (i && rows[i] < rows[i - 1])) return fail(cudaErrorInvalidValue);
            rowptr[rows[i] + 1]++;
        }
        for (uint32_t r = 0; r < n_rows; r++) rowptr[r + 1] += rowptr[r];
    } else if (part_rows_dev) { // column-direction chunks: part i lies in row part_rows[i]
        std::vector<uint32_t> pr(n_parts);
        if ((e = cudaMemcpyAsync(pr.data(), part_rows_dev, (size_t)n_parts * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = pr[n_parts - 1] + 1;
        rowptr.assign((size_t)n_rows + 1, 0);
        for (uint32_t i = 0; i < n_parts; i++) {
            if (pr[i] >= n_rows || (i && pr[i] < pr[i - 1]) || parts[i + 1] < parts[i]) return fail(cudaErrorInvalidValue);
            rowptr[pr[i] + 1] += parts[i + 1] - parts[i];
        }
        for (uint32_t r = 0; r < n_rows; r++) rowptr[r + 1] += rowptr[r];
    } else if (parts_are_rows) {
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = n_parts, rowptr = parts;
    } else return fail(cudaErrorInvalidValue);
    if (parts[0] != 0 || parts[n_parts] > nnz || rowptr[n_rows] < parts[n_parts]) return fail(cudaErrorInvalidValue);

    // segments: the chain's parts, long ones cut into equal slices (GS_MERGE_CAP non-zeros at most)
    const bool per_thread = (uint64_t)p.N * (p.a.val_is_f16 ? 2 : 4) <= (uint64_t)env_int("GS_MERGE_CHUNK_ROW_BYTES", 256) && env_int("GS_MERGE_THREAD", 1) != 0;
    const int LG = per_thread ? (p.a.val_is_f16 ? chunk_lanes<__half>(p.N) : chunk_lanes<float>(p.N)) : p.N < 32 ? p.N : 32;
    const uint32_t cap = (uint32_t)std::max(LG, env_int("GS_MERGE_CAP", per_thread ? (LG >= 4 ? 64 : LG == 2 ? 40 : 32) : LG == 32 ? 512 : 32 * LG)); // chunk kernel: every lane of a group walks the whole segment
    pl->per_thread = per_thread, pl->cap = cap;
    if (per_thread) { // staging needs 16-byte aligned streams and at most 48 KB for a CTA's piece
        const uint32_t elems = (128 / LG) * cap + 8;
        const uint64_t bytes = 4ull * (stage_words(elems, p.a.col_is_u16 ? 2 : 4) + stage_words(elems, p.a.val_is_f16 ? 2 : 4));
        pl->stage = env_int("GS_MERGE_STAGE", LG <= 4) != 0 && LG <= 8 && bytes <= 48 * 1024 && (((uintptr_t)p.a.col | (uintptr_t)p.a.val) & 15) == 0;
    }
    std::vector<uint32_t> seg_nz, seg_row;
    for (uint32_t i = 0; i < n_parts; i++) {
        const uint32_t a = parts[i], b = parts[i + 1];
        if (b < a) return fail(cudaErrorInvalidValue);
        if (b == a) continue;
        const uint32_t pieces = (b - a + cap - 1) / cap;
        for (uint32_t q = 0; q < pieces; q++) {
            const uint32_t lo = a + (uint32_t)((uint64_t)(b - a) * q / pieces);
            seg_nz.push_back(lo);
            seg_row.push_back((uint32_t)(std::upper_bound(rowptr.begin(), rowptr.end(), lo) - rowptr.begin()) - 1);
        }
    }
    seg_nz.push_back(parts[n_parts]);
    pl->n_seg = (uint32_t)seg_row.size(), pl->n_rows = n_rows, pl->N = p.N, pl->part_bytes = ((uint64_t)n_parts + 1) * 4;
    if (pl->n_seg == 0) return fail(cudaErrorInvalidValue);
    auto up = [&](uint32_t **dst, const std::vector<uint32_t> &v) {
        cudaError_t e2 = cudaMalloc((void **)dst, v.size() * 4 + 16);
        return e2 != cudaSuccess ? e2 : cudaMemcpyAsync(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice, s);
    };
    if ((e = up(&pl->seg_nz, seg_nz)) != cudaSuccess || (e = up(&pl->seg_row, seg_row)) != cudaSuccess || (e = up(&pl->rowptr, rowptr)) != cudaSuccess) return fail(e);
    // ticket mode: who shares which row
    // tickets by default where they are needed (HALF: red.add would round every slice's partial to fp16) or measured at least as fast as memset +
    // red.add (rows of B up to 128 bytes); wider fp32 rows keep red.add: the partial buffer costs 8 N bytes per segment there (C5 uniform, N = 128:
    // 1 752 us with tickets against 1 493 with red.add)
    const bool ticket_default = p.a.val_is_f16 || (per_thread && (uint64_t)p.N * 4 <= 128);
    if (p.a.begin_row == 0 && env_int("GS_MERGE_TICKET", ticket_default ? 1 : 0) != 0) { // (a plan on a sub-matrix keeps the memset of the whole of C)
        const uint32_t n_seg = pl->n_seg;
        std::vector<uint32_t> head(n_seg), tail(n_seg), zero;
        auto seg_of = [&](uint32_t nz) { // the segment that holds stream position nz
            const uint32_t q = (uint32_t)(std::upper_bound(seg_nz.begin(), seg_nz.end(), nz) - seg_nz.begin());
            return std::min(q ? q - 1 : 0u, n_seg - 1);
        };
        uint32_t most = 1;
        for (uint32_t q = 0; q < n_seg; q++) {
            const uint32_t a = seg_nz[q], b = seg_nz[q + 1], f = seg_row[q];
            const uint32_t l = (uint32_t)(std::upper_bound(rowptr.begin(), rowptr.end(), b - 1) - rowptr.begin()) - 1;
            head[q] = rowptr[f] < a ? seg_of(rowptr[f]) : q;
            tail[q] = rowptr[l + 1] > b ? seg_of(rowptr[l + 1] - 1) : q;
            most = std::max(most, std::max(q - head[q], tail[q] - q) + 1);
        }
        for (uint32_t r = 0; r < n_rows; r++)
            if (rowptr[r + 1] == rowptr[r] || rowptr[r] >= parts[n_parts]) zero.push_back(r);
        for (uint32_t r = n_rows; r < p.a.rows; r++) zero.push_back(r);
        pl->max_sharers = most;
        // (the partial sums take 8 N bytes per segment: beyond GS_MERGE_TICKET_MB the plan stays with red.add)
        if (most <= (uint32_t)env_int("GS_MERGE_TICKET_MAX", 64) && (uint64_t)n_seg * 2 * p.N * sizeof(float) <= (uint64_t)env_int("GS_MERGE_TICKET_MB", 512) << 20) {
            if ((e = up(&pl->head_s0, head)) != cudaSuccess || (e = up(&pl->tail_s1, tail)) != cudaSuccess) return fail(e);
            if (!zero.empty() && (e = up(&pl->zero_rows, zero)) != cudaSuccess) return fail(e);
            pl->n_zero = (uint32_t)zero.size();
            if ((e = cudaMalloc((void **)&pl->ticket, (size_t)n_seg * 4)) != cudaSuccess || (e = cudaMemsetAsync(pl->ticket, 0, (size_t)n_seg * 4, s)) != cudaSuccess) return fail(e);
            if ((e = cudaMalloc((void **)&pl->partial, (size_t)n_seg * 2 * p.N * sizeof(float))) != cudaSuccess) return fail(e);
        }
    }
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
    snprintf(pl->name, sizeof pl->name, "k_b200_merge%s<N%d,%s,%s,%d lane%s/segment,%s%s%u segments of <=%u nnz from %u parts>", per_thread ? "_chunk" : "", p.N, p.a.col_is_u16 ? "u16" : "u32",
             p.a.val_is_f16 ? "f16" : "f32", LG, LG > 1 ? "s" : "", pl->stage ? "staged," : "", pl->ticket ? "tickets," : "", pl->n_seg, cap, n_parts);
    *out = pl;
    return cudaSuccess;
}

cudaError_t b200_merge_launch(const b200_merge_plan *pl, const launch_params &p, cudaStream_t s)
{
    merge_args a;
    a.col = p.a.col, a.val = p.a.val, a.B = p.B, a.C = p.C, a.seg_nz = pl->seg_nz, a.seg_row = pl->seg_row, a.rowptr = pl->rowptr, a.perm = p.a.perm;
    a.n_seg = pl->n_seg, a.begin_row = p.a.begin_row, a.N = pl->N;
    a.nnz = (uint32_t)p.a.nnz, a.stage_col_words = 0;
    a.head_s0 = pl->head_s0, a.tail_s1 = pl->tail_s1, a.ticket = pl->ticket, a.partial = pl->partial, a.zero_rows = pl->zero_rows, a.n_zero = pl->n_zero;
    if (pl->per_thread) {
        const int lanes = p.a.val_is_f16 ? chunk_lanes<__half>(pl->N) : chunk_lanes<float>(pl->N);
        // the plan's decision holds for the arrays it was made on; a stream that moved to an unaligned address runs unstaged
        const bool stage = pl->stage && (((uintptr_t)p.a.col | (uintptr_t)p.a.val) & 15) == 0;
        const bool tk = pl->ticket != nullptr;
        const void *tf = p.a.col_is_u16 ? (p.a.val_is_f16 ? pick_thread<uint16_t, __half>(pl->N, stage, tk) : pick_thread<uint16_t, float>(pl->N, stage, tk))
                                        : (p.a.val_is_f16 ? pick_thread<uint32_t, __half>(pl->N, stage, tk) : pick_thread<uint32_t, float>(pl->N, stage, tk));
        if (!tf) return cudaErrorInvalidValue;
        void *targs[] = {&a};
        const uint32_t elems = (128 / lanes) * pl->cap + 8;
        a.stage_col_words = stage_words(elems, p.a.col_is_u16 ? 2 : 4);
        const size_t smem = stage ? 4ull * (a.stage_col_words + stage_words(elems, p.a.val_is_f16 ? 2 : 4)) : 0;
        const uint64_t threads = (uint64_t)pl->n_seg * lanes;
        // programmatic dependent launch whenever the engine allows it: the kernel touches nothing but the plan's tables and the stream of A
        // before griddepcontrol.wait, so it does not matter who ran before it.  Not behind the memset of the red.add mode (a memset node
        // is not a kernel: the attribute would be ignored, and the order must hold).
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3((unsigned)((threads + 127) / 128)), cfg.blockDim = dim3(128), cfg.dynamicSmemBytes = smem, cfg.stream = s;
        cfg.attrs = attr, cfg.numAttrs = (p.pdl && pl->ticket) ? 1 : 0;
        return cudaLaunchKernelExC(&cfg, tf, targs);
    }
    const bool tk = pl->ticket != nullptr;
    const void *fn = p.a.col_is_u16 ? (p.a.val_is_f16 ? pick<uint16_t, __half>(pl->N, tk) : pick<uint16_t, float>(pl->N, tk))
                                    : (p.a.val_is_f16 ? pick<uint32_t, __half>(pl->N, tk) : pick<uint32_t, float>(pl->N, tk));
    if (!fn) return cudaErrorInvalidValue;
    const int LG = pl->N < 32 ? pl->N : 32;
    const uint64_t threads = (uint64_t)pl->n_seg * LG;
    const unsigned grid = (unsigned)((threads + MERGE_WARPS * 32 - 1) / (MERGE_WARPS * 32));
    void *args[] = {&a};
    return cudaLaunchKernel(fn, dim3(grid), dim3(MERGE_WARPS * 32), args, 0, s);
}

void b200_merge_free(b200_merge_plan *pl)
{
    if (!pl) return;
    cudaFree(pl->seg_nz), cudaFree(pl->seg_row), cudaFree(pl->rowptr);
    cudaFree(pl->head_s0), cudaFree(pl->tail_s1), cudaFree(pl->ticket), cudaFree(pl->zero_rows), cudaFree(pl->partial);
    delete pl;
}

const char *b200_merge_kernel_name(const b200_merge_plan *pl) { return pl ? pl->name : ""; }

// red.add mode: C must be zero when the kernel starts; ticket mode writes every row itself
bool b200_merge_needs_zero(const b200_merge_plan *pl) { return !pl || !pl->ticket; }

// what one launch streams: A once, the segment table and the row pointer, B and C once each
uint64_t b200_merge_bytes(const b200_merge_plan *pl, uint64_t nnz, uint64_t idx_bytes, uint64_t val_bytes, uint64_t rows, uint64_t cols)
{
    return nnz * (idx_bytes + val_bytes) + ((uint64_t)pl->n_seg * 2 + 1 + pl->n_rows + 1) * 4 + (cols + rows) * pl->N * val_bytes;
}

} // namespace gsd
