// Device code of the B200 stream schedule of the row kernel (host side and design notes: kernels_b200.cu).
//
// v2 (one warp per row) measured 23 % of the launch with the SMs idle (launch + tail), long_scoreboard as the top
// stall and the shared-memory data pipe -- not HBM -- as the busiest unit (profiles/r1_row_v2_summary.md).  This
// schedule keeps v2's inner step (8 non-zeros per lane, 128-bit loads, FHFMA, swizzled panel) and changes the
// work split:
//   * the non-zero stream of a CTA (whole rows, nnz-balanced over the CTAs) is cut into 256-wide steps and the steps
//     are dealt evenly to the warps: every warp runs the same number of steps whatever the row lengths are;
//   * a row that crosses a warp boundary is finished through shared memory (its owner adds the partial sums of
//     the following warps after one __syncthreads); CTAs own whole rows, so C needs no atomics and no memset;
//   * D steps of A per warp are in flight in registers, across row boundaries;
//   * the CTA's row table (rowptr slice, output row ids) and its rows of C live in shared memory, so the step loop
//     issues no long-latency load but the A stream and C is written once, coalesced, at the end;
//   * programmatic dependent launch: the next launch's CTAs become resident as soon as an SM has room, stage their
//     panel and run their whole step loop while this grid drains; the only global stores of the kernel -- the C
//     tile -- come after griddepcontrol.wait.
#pragma once
#include "b200_row.cuh"

namespace gsd {
namespace row {

struct stream_args {
    const void *col, *val, *B;
    void *C;
    const uint32_t *rowptr, *rowid, *perm;
    const uint32_t *w_nz;  // grid x (W+1): first non-zero of each warp's range; entry W = end of the CTA's range
    const uint32_t *w_row; // grid x (W+1): first row that STARTS at or after w_nz; entry W = the CTA's row end
    uint32_t n_rows, begin_row, b_bytes, bulk_panel;
    uint32_t pf; // > 0: also prefetch the A lines of step (D + pf) into L2
};

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename ColT, typename ValT> struct a_chunk {
    uint4 c[sizeof(ColT) / 2]; // 8 column ids
    uint4 v[sizeof(ValT) / 2]; // 8 values
    __device__ __forceinline__ void load(const ColT *col, const ValT *val, uint32_t pos, uint32_t end)
    {
        if (pos < end) {
            c[0] = ld_stream_v4(col + pos);
            if (sizeof(ColT) == 4) c[sizeof(ColT) / 2 - 1] = ld_stream_v4(col + pos + 4);
            v[0] = ld_stream_v4(val + pos);
            if (sizeof(ValT) == 4) v[sizeof(ValT) / 2 - 1] = ld_stream_v4(val + pos + 4);
        }
    }
};

__device__ __forceinline__ void to_val8(const a_chunk<uint16_t, __half> &ch, val8<__half> &v) { v.r = ch.v[0]; }
__device__ __forceinline__ void to_val8(const a_chunk<uint32_t, __half> &ch, val8<__half> &v) { v.r = ch.v[0]; }
__device__ __forceinline__ void to_val8(const a_chunk<uint16_t, float> &ch, val8<float> &v) { v.r0 = ch.v[0], v.r1 = ch.v[1]; }
__device__ __forceinline__ void to_val8(const a_chunk<uint32_t, float> &ch, val8<float> &v) { v.r0 = ch.v[0], v.r1 = ch.v[1]; }

// Sum every accumulator over the 32 lanes with NT-1 + (5 - log2 NT) shuffles (a transposing butterfly: each stage
// halves the number of live accumulators per lane).  Afterwards lane l holds the total of dense column
// l >> (5 - log2 NT); the value is replicated over the 32/NT lanes that share a column.
template <int NT> __device__ __forceinline__ float warp_reduce_cols(acc_t<NT> &acc, int lane)
{
    static_assert(NT == 8 || NT == 16 || NT == 32, "dense widths of the shared-memory panel kernels");
    int m = 16;
#pragma unroll
    for (int n = NT / 2; n >= 1; n >>= 1) {
        const bool up = (lane & m) != 0;
#pragma unroll
        for (int i = 0; i < n; i++) {
            float send = up ? acc.v[i] : acc.v[i + n];
            float keep = up ? acc.v[i + n] : acc.v[i];
            acc.v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
        }
        m >>= 1;
    }
    float s = acc.v[0];
#pragma unroll
    for (int k = 32 / NT / 2; k >= 1; k >>= 1) s += __shfl_xor_sync(0xffffffffu, s, k);
    return s;
}

__device__ __forceinline__ void store_c(__half *C, size_t idx, float s) { C[idx] = __float2half_rn(s); }
__device__ __forceinline__ void store_c(float *C, size_t idx, float s) { C[idx] = s; }

// One step of a lane: 8 non-zeros against the panel.  PRE (u16 ids, fp16, one 16-byte piece per dense row and
// K * 16 <= 65536): the layout pass stored id * 16, i.e. the byte offset of the dense row inside the panel, so an
// address is one mask or shift and the panel base folds into the LDS immediate.
template <int NT, bool SWZ, bool PRE, typename ColT, typename ValT> struct gather8 {
    static __device__ __forceinline__ void run(acc_t<NT> &acc, const a_chunk<ColT, ValT> &ch, const unsigned char *panel)
    {
        uint32_t c[8];
        unpack_cols<ColT>(ch.c[0], ch.c[sizeof(ColT) / 2 - 1], c);
        val8<ValT> v;
        to_val8(ch, v);
        step8<NT, SWZ>(acc, c, v, panel);
    }
};
template <> struct gather8<8, false, true, uint16_t, __half> {
    static __device__ __forceinline__ void run(acc_t<8> &acc, const a_chunk<uint16_t, __half> &ch, const unsigned char *panel)
    {
        const uint32_t cw[4] = {ch.c[0].x, ch.c[0].y, ch.c[0].z, ch.c[0].w}, vw[4] = {ch.v[0].x, ch.v[0].y, ch.v[0].z, ch.v[0].w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint4 b0 = *(const uint4 *)(panel + (cw[j] & 0xffffu)), b1 = *(const uint4 *)(panel + (cw[j] >> 16));
            unsigned short a0, a1, h0, h1;
            split_h2(vw[j], a0, a1);
            split_h2(b0.x, h0, h1), fma_h(acc.v[0], a0, h0), fma_h(acc.v[1], a0, h1);
            split_h2(b0.y, h0, h1), fma_h(acc.v[2], a0, h0), fma_h(acc.v[3], a0, h1);
            split_h2(b0.z, h0, h1), fma_h(acc.v[4], a0, h0), fma_h(acc.v[5], a0, h1);
            split_h2(b0.w, h0, h1), fma_h(acc.v[6], a0, h0), fma_h(acc.v[7], a0, h1);
            split_h2(b1.x, h0, h1), fma_h(acc.v[0], a1, h0), fma_h(acc.v[1], a1, h1);
            split_h2(b1.y, h0, h1), fma_h(acc.v[2], a1, h0), fma_h(acc.v[3], a1, h1);
            split_h2(b1.z, h0, h1), fma_h(acc.v[4], a1, h0), fma_h(acc.v[5], a1, h1);
            split_h2(b1.w, h0, h1), fma_h(acc.v[6], a1, h0), fma_h(acc.v[7], a1, h1);
        }
    }
};

// NT = dense width, D = steps of A in flight per warp, SHAPE = resident CTAs per SM the launch bounds allow
// (1: one CTA of <= 32 warps, 2: two CTAs of <= 16 warps; 64 registers per thread either way).
// Dynamic shared memory: the dense panel, then the CTA's row table (rowptr slice + output row of every row) so
// that the only long-latency loads inside the step loop are the A stream itself.
template <int NT, typename ColT, typename ValT, int D, int SHAPE, bool PRE>
__global__ void __launch_bounds__(SHAPE == 1 ? 1024 : SHAPE == 2 ? 512 : 320, SHAPE) k_b200_stream(stream_args a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ float head_val[32][NT];
    __shared__ uint32_t head_flag[32]; // 0: no head row, 1: head row ended inside the warp, 2: the whole range lies inside it
    constexpr int ROWB = NT * (int)sizeof(ValT);
    constexpr bool SWZ = ROWB > 16 && ROWB <= 128;
    constexpr int REP = 32 / NT; // lanes that share a dense column after warp_reduce_cols
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val;
    ValT *C = (ValT *)a.C;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, W = blockDim.x >> 5;
    const int my_col = lane / REP;
    const bool writer = (lane % REP) == 0;

    const uint32_t cta0 = blockIdx.x * (W + 1);
    const uint32_t nb = a.w_nz[cta0 + warp], ne = a.w_nz[cta0 + warp + 1];
    const uint32_t r0 = a.w_row[cta0], r_own = a.w_row[cta0 + warp], r_end = a.w_row[cta0 + W];
    uint32_t *s_rp = (uint32_t *)(smem + a.b_bytes); // rowptr[r0 .. r_end]
    uint32_t *s_out = s_rp + (r_end - r0 + 1);       // output row of r0 .. r_end - 1
    float *s_c = (float *)(s_out + (r_end - r0));    // the CTA's rows of C (fp32), stored once at the very end

    const bool bulk = !SWZ && a.bulk_panel;
    if (bulk && threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_expect_tx(&bar, a.b_bytes);
        for (uint32_t off = 0; off < a.b_bytes; off += 32768u) {
            uint32_t n = a.b_bytes - off < 32768u ? a.b_bytes - off : 32768u;
            bulk_g2s(smem + off, (const unsigned char *)a.B + off, n, &bar);
        }
    }

    // the first D steps of A go out before the panel has landed
    a_chunk<ColT, ValT> q[D];
#pragma unroll
    for (int s = 0; s < D; s++) q[s].load(col, val, nb + s * 256 + lane * 8, ne);

    for (uint32_t i = threadIdx.x; i <= r_end - r0; i += blockDim.x) {
        s_rp[i] = a.rowptr[r0 + i];
        if (i < r_end - r0) {
            uint32_t row = a.rowid[r0 + i];
            s_out[i] = (a.perm ? a.perm[row] : row) + a.begin_row;
        }
    }
    for (uint32_t i = threadIdx.x; i < (r_end - r0) * NT; i += blockDim.x) s_c[i] = 0.0f; // rows without non-zeros stay 0
    if (lane == 0) head_flag[warp] = 0;
    if (!bulk) {
        constexpr int P = ROWB / 16;
        const uint4 *src = (const uint4 *)a.B;
        uint32_t n_pieces = a.b_bytes >> 4;
        for (uint32_t x = threadIdx.x; x < n_pieces; x += blockDim.x) *(uint4 *)(smem + piece_offset<ROWB, SWZ>(x / P, x % P)) = __ldg(src + x);
    }
    __syncthreads();
    if (bulk) mbar_wait(&bar, 0);
    pdl_launch_dependents();

    // row state: `cur` is the row the next non-zero belongs to, `hi` its end
    uint32_t cur = r_own;
    bool is_head = false;
    if (nb < ne && s_rp[r_own - r0] > nb) cur = r_own - 1, is_head = true; // the range starts inside row r_own - 1
    uint32_t hi = cur < r_end ? s_rp[cur + 1 - r0] : ne;

    acc_t<NT> acc;
#pragma unroll
    for (int i = 0; i < NT; i++) acc.v[i] = 0.0f;

    // the row `cur` is complete: reduce over the lanes; a head row goes to its owner through head_val, an owned row
    // to the CTA's C tile (nothing touches global C before griddepcontrol.wait at the very end)
    auto finish = [&]() {
        float s = warp_reduce_cols<NT>(acc, lane);
#pragma unroll
        for (int i = 0; i < NT; i++) acc.v[i] = 0.0f;
        if (is_head) {
            if (writer) head_val[warp][my_col] = s;
            if (lane == 0) head_flag[warp] = 1;
        } else if (writer) {
            s_c[(cur - r0) * NT + my_col] = s;
        }
    };

    for (uint32_t base = nb; base < ne; base += 256u * D) {
#pragma unroll
        for (int s = 0; s < D; s++) {
            const uint32_t b = base + s * 256u;
            if (b >= ne) break;
            const a_chunk<ColT, ValT> ch = q[s];
            q[s].load(col, val, b + 256u * D + lane * 8, ne);
            if (a.pf) {
                const uint32_t pp = b + 256u * (D + a.pf) + lane * 8;
                if (pp < ne && (lane & 3) == 0) {
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(col + pp));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(val + pp));
                }
            }
            const uint32_t pos = b + lane * 8;
            const uint32_t step_end = b + 256u < ne ? b + 256u : ne;
            bool live = pos < step_end;
            for (;;) {
                if (live && pos < hi) gather8<NT, SWZ, PRE, ColT, ValT>::run(acc, ch, smem);
                if (hi > step_end || cur >= r_end) break; // the row goes on in the next step
                finish();
                live = live && pos >= hi;
                const uint32_t next_start = hi;
                cur++, is_head = false;
                if (cur >= r_end) break;
                hi = s_rp[cur + 1 - r0];
                if (next_start >= step_end) break;
            }
        }
    }

    // a row still open at the end of the range: either this warp owns it (it started here) or the whole range lies inside it
    float tail = 0.0f;
    bool own_tail = false;
    if (nb < ne && cur < r_end && (is_head || s_rp[cur - r0] < ne)) {
        tail = warp_reduce_cols<NT>(acc, lane);
        if (is_head) {
            if (writer) head_val[warp][my_col] = tail;
            if (lane == 0) head_flag[warp] = 2;
        } else {
            own_tail = true;
        }
    }
    __syncthreads();
    if (own_tail) {
        for (int w2 = warp + 1; w2 < W; w2++) {
            uint32_t f = head_flag[w2];
            if (!f) break;
            tail += head_val[w2][my_col];
            if (f == 1) break;
        }
        if (writer) s_c[(cur - r0) * NT + my_col] = tail;
    }
    __syncthreads();
    // the CTA's rows of C: the only global stores of the kernel, after the previous grid on the stream has completed
    pdl_wait();
    for (uint32_t i = threadIdx.x; i < (r_end - r0) * NT; i += blockDim.x) store_c(C, (size_t)s_out[i / NT] * NT + (i % NT), s_c[i]);
}

// ---- B200 layout pass for the stream schedule ------------------------------------------------------------------------
// Same idea as k_rainbow_rows (b200_row.cuh): inside a row the non-zeros may be stored in any order, so they are
// placed such that the 8 lanes of a quarter warp -- which share one 128-bit LDS wavefront -- gather from 8 distinct
// bank groups (col % 8).  Here the 256-wide steps are laid from the first non-zero of the row's CTA, not from the row
// start, so a row covers a run of lane slots [a, b) of that grid: a partial first quarter, whole quarters, a partial
// last quarter.  Lanes of other rows in the same quarter run in a separate masked pass and do not conflict.
// The row is bucket-sorted by col % 8 and dealt in 8 rounds; round t fills lane t of every quarter that has one.
template <typename ColT, typename ValT>
__global__ void k_rainbow_stream(const ColT *col, const ValT *val, ColT *col2, ValT *val2, ColT *tc, ValT *tv, const uint32_t *rowptr, uint32_t n_rows,
                                 const uint32_t *cta_row, uint32_t n_cta, uint32_t scale)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    uint32_t lo = rowptr[r], hi = rowptr[r + 1], L = hi - lo;
    if (L == 0) return;
    uint32_t c0 = 0, c1 = n_cta; // last CTA whose first row is <= r
    while (c1 - c0 > 1) {
        uint32_t mid = (c0 + c1) / 2;
        if (cta_row[mid] <= r) c0 = mid;
        else c1 = mid;
    }
    const uint32_t nz0 = rowptr[cta_row[c0]];

    uint32_t cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0}, start[8];
    for (uint32_t i = 0; i < L; i++) cnt[col[lo + i] & 7]++;
    uint32_t s = 0;
    for (int k = 0; k < 8; k++) start[k] = s, s += cnt[k];
    for (uint32_t i = 0; i < L; i++) {
        uint32_t k = start[col[lo + i] & 7]++;
        tc[lo + k] = col[lo + i], tv[lo + k] = val[lo + i];
    }

    const uint32_t a = (lo - nz0) / 8, b = (hi - nz0) / 8; // lane slots of the CTA's step grid
    const uint32_t qa = a / 8, qb = (b + 7) / 8, nq = qb - qa;
    const uint32_t m_first = (b < (qa + 1) * 8 ? b : (qa + 1) * 8) - a;
    const uint32_t m_last = nq > 1 ? b - (qb - 1) * 8 : 0;
    uint32_t k = 0;
    for (uint32_t t = 0; t < 8; t++) {
        for (uint32_t i = 0; i < nq; i++) {
            const uint32_t cap = i == 0 ? m_first : (i == nq - 1 ? m_last : 8u);
            if (cap <= t) continue;
            const uint32_t slot0 = i == 0 ? a : (qa + i) * 8;
            const uint32_t p = nz0 + (slot0 + t) * 8;
            for (uint32_t j = 0; j < 8; j++, k++) col2[p + j] = (ColT)(tc[lo + k] * scale), val2[p + j] = tv[lo + k]; // scale 16: PRE ids
        }
    }
}

} // namespace row
} // namespace gsd
