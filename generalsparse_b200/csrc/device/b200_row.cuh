// Device code of the B200 row kernel (see kernels_b200.cu for the design notes and the host side).
#pragma once
#include "common.cuh"

namespace gsd {
namespace row {

struct row_args {
    const void *col, *val, *B;
    void *C;
    const uint32_t *rowptr, *rowid, *perm;
    uint32_t n_rows, begin_row, K;
    uint32_t b_bytes; // bytes of the dense panel
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(
                     smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}

// acc += a * b with fp16 inputs and an fp32 accumulator, one rounding (SASS: FHFMA)
__device__ __forceinline__ void fma_h(float &acc, unsigned short a, unsigned short b) { asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(acc) : "h"(a), "h"(b)); }
__device__ __forceinline__ void split_h2(uint32_t x, unsigned short &lo, unsigned short &hi) { asm("mov.b32 {%0,%1}, %2;" : "=h"(lo), "=h"(hi) : "r"(x)); }

template <int NT> struct acc_t {
    float v[NT];
};

// Piece i (16 bytes) of dense row c.  A row is P = ROWB/16 pieces; with SWZ piece i lives at slot
// i ^ ((c / (8/P)) % P) -- the 32/64/128-byte TMA swizzle patterns -- so that for any fixed i the bank group a
// lane touches is c % 8: lanes whose column ids differ mod 8 never collide, whatever the dense width.
template <int ROWB, bool SWZ> __device__ __forceinline__ uint32_t piece_offset(uint32_t c, int i)
{
    constexpr int P = ROWB / 16;
    if (SWZ && P > 1 && P <= 8) {
        constexpr int Q = 8 / P;
        return c * ROWB + ((i ^ ((c / Q) % P)) << 4);
    }
    return c * ROWB + (i << 4);
}

template <int NT, bool SWZ> __device__ __forceinline__ void nz_fma(acc_t<NT> &acc, unsigned short a, const unsigned char *panel, uint32_t c, __half)
{
#pragma unroll
    for (int q = 0; q < NT / 8; q++) {
        uint4 b = *(const uint4 *)(panel + piece_offset<NT * 2, SWZ>(c, q));
        unsigned short h0, h1;
        split_h2(b.x, h0, h1), fma_h(acc.v[q * 8 + 0], a, h0), fma_h(acc.v[q * 8 + 1], a, h1);
        split_h2(b.y, h0, h1), fma_h(acc.v[q * 8 + 2], a, h0), fma_h(acc.v[q * 8 + 3], a, h1);
        split_h2(b.z, h0, h1), fma_h(acc.v[q * 8 + 4], a, h0), fma_h(acc.v[q * 8 + 5], a, h1);
        split_h2(b.w, h0, h1), fma_h(acc.v[q * 8 + 6], a, h0), fma_h(acc.v[q * 8 + 7], a, h1);
    }
}
template <int NT, bool SWZ> __device__ __forceinline__ void nz_fma(acc_t<NT> &acc, float a, const unsigned char *panel, uint32_t c, float)
{
#pragma unroll
    for (int q = 0; q < NT / 4; q++) {
        float4 b = *(const float4 *)(panel + piece_offset<NT * 4, SWZ>(c, q));
        acc.v[q * 4 + 0] = fmaf(a, b.x, acc.v[q * 4 + 0]);
        acc.v[q * 4 + 1] = fmaf(a, b.y, acc.v[q * 4 + 1]);
        acc.v[q * 4 + 2] = fmaf(a, b.z, acc.v[q * 4 + 2]);
        acc.v[q * 4 + 3] = fmaf(a, b.w, acc.v[q * 4 + 3]);
    }
}

template <typename ColT> __device__ __forceinline__ void load_cols8(const ColT *p, uint32_t (&c)[8]);
template <> __device__ __forceinline__ void load_cols8<uint16_t>(const uint16_t *p, uint32_t (&c)[8])
{
    uint4 r = ld_stream_v4(p);
    c[0] = r.x & 0xffffu, c[1] = r.x >> 16, c[2] = r.y & 0xffffu, c[3] = r.y >> 16;
    c[4] = r.z & 0xffffu, c[5] = r.z >> 16, c[6] = r.w & 0xffffu, c[7] = r.w >> 16;
}
template <> __device__ __forceinline__ void load_cols8<uint32_t>(const uint32_t *p, uint32_t (&c)[8])
{
    uint4 r0 = ld_stream_v4(p), r1 = ld_stream_v4(p + 4);
    c[0] = r0.x, c[1] = r0.y, c[2] = r0.z, c[3] = r0.w, c[4] = r1.x, c[5] = r1.y, c[6] = r1.z, c[7] = r1.w;
}

template <typename ValT> struct val8;
template <> struct val8<__half> {
    uint4 r;
    __device__ __forceinline__ void load(const __half *p) { r = ld_stream_v4(p); }
};
template <> struct val8<float> {
    uint4 r0, r1;
    __device__ __forceinline__ void load(const float *p) { r0 = ld_stream_v4(p), r1 = ld_stream_v4(p + 4); }
};

template <int NT, bool SWZ> __device__ __forceinline__ void step8(acc_t<NT> &acc, const uint32_t (&c)[8], const val8<__half> &v, const unsigned char *panel)
{
    const uint32_t w[4] = {v.r.x, v.r.y, v.r.z, v.r.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
        unsigned short a0, a1;
        split_h2(w[j], a0, a1);
        nz_fma<NT, SWZ>(acc, a0, panel, c[2 * j], __half());
        nz_fma<NT, SWZ>(acc, a1, panel, c[2 * j + 1], __half());
    }
}
template <int NT, bool SWZ> __device__ __forceinline__ void step8(acc_t<NT> &acc, const uint32_t (&c)[8], const val8<float> &v, const unsigned char *panel)
{
    const uint32_t w[8] = {v.r0.x, v.r0.y, v.r0.z, v.r0.w, v.r1.x, v.r1.y, v.r1.z, v.r1.w};
#pragma unroll
    for (int j = 0; j < 8; j++) nz_fma<NT, SWZ>(acc, __uint_as_float(w[j]), panel, c[j], float());
}

template <int NT, bool SWZ> __device__ __forceinline__ void nz1(acc_t<NT> &acc, __half v, uint32_t c, const unsigned char *panel)
{
    nz_fma<NT, SWZ>(acc, __half_as_ushort(v), panel, c, __half());
}
template <int NT, bool SWZ> __device__ __forceinline__ void nz1(acc_t<NT> &acc, float v, uint32_t c, const unsigned char *panel)
{
    nz_fma<NT, SWZ>(acc, v, panel, c, float());
}

template <int NT> __device__ __forceinline__ void store_out(__half *C, size_t row, const acc_t<NT> &acc)
{
    uint4 *dst = (uint4 *)(C + row * NT);
#pragma unroll
    for (int q = 0; q < NT / 8; q++) {
        __half2 h0 = __floats2half2_rn(acc.v[q * 8 + 0], acc.v[q * 8 + 1]), h1 = __floats2half2_rn(acc.v[q * 8 + 2], acc.v[q * 8 + 3]);
        __half2 h2 = __floats2half2_rn(acc.v[q * 8 + 4], acc.v[q * 8 + 5]), h3 = __floats2half2_rn(acc.v[q * 8 + 6], acc.v[q * 8 + 7]);
        uint4 o;
        o.x = *(uint32_t *)&h0, o.y = *(uint32_t *)&h1, o.z = *(uint32_t *)&h2, o.w = *(uint32_t *)&h3;
        dst[q] = o;
    }
}
template <int NT> __device__ __forceinline__ void store_out(float *C, size_t row, const acc_t<NT> &acc)
{
    float4 *dst = (float4 *)(C + row * NT);
#pragma unroll
    for (int q = 0; q < NT / 4; q++) dst[q] = make_float4(acc.v[q * 4], acc.v[q * 4 + 1], acc.v[q * 4 + 2], acc.v[q * 4 + 3]);
}

// NT = N (dense width); VEC = 8 (every row offset a multiple of 8) or 1; SMEM = panel staged in shared memory.
// One warp per row, rows taken with stride = number of warps in the grid.
template <int NT, typename ColT, typename ValT, int VEC, bool SMEM> __global__ void __launch_bounds__(1024, 1) k_b200_row(row_args a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    constexpr int ROWB = NT * (int)sizeof(ValT);
    constexpr bool SWZ = SMEM && ROWB > 16 && ROWB <= 128;
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val;
    const unsigned char *panel = (const unsigned char *)a.B;
    const int lane = threadIdx.x & 31;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const uint32_t n_warps = gridDim.x * warps_per_cta;
    // interleave the CTAs' warps so that consecutive rows land on different SMs
    uint32_t r = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;

    if (SMEM && !SWZ) {
        if (threadIdx.x == 0) {
            mbar_init(&bar, 1);
            mbar_expect_tx(&bar, a.b_bytes);
            for (uint32_t off = 0; off < a.b_bytes; off += 32768u) { // one bulk copy per 32 KiB piece, one barrier
                uint32_t n = a.b_bytes - off < 32768u ? a.b_bytes - off : 32768u;
                bulk_g2s(smem + off, (const unsigned char *)a.B + off, n, &bar);
            }
        }
        panel = smem;
    }

    uint32_t lo = 0, hi = 0;
    if (r < a.n_rows) lo = a.rowptr[r], hi = a.rowptr[r + 1];
    uint32_t c_next[8];
    val8<ValT> v_next;
    bool have_next = false;
    if (VEC == 8) { // first A loads go out before the panel has landed
        uint32_t nz = lo + lane * 8;
        if (nz < hi) {
            load_cols8<ColT>(col + nz, c_next);
            v_next.load(val + nz);
            have_next = true;
        }
    }
    if (SMEM && SWZ) {
        // swizzled staging: 16-byte pieces, coalesced global reads, XOR-permuted inside each 128-byte line
        constexpr int P = ROWB / 16;
        const uint4 *src = (const uint4 *)a.B;
        uint32_t n_pieces = a.b_bytes >> 4;
        for (uint32_t x = threadIdx.x; x < n_pieces; x += blockDim.x) {
            uint32_t c = x / P;
            int i = x % P;
            *(uint4 *)(smem + piece_offset<ROWB, true>(c, i)) = __ldg(src + x);
        }
        panel = smem;
        __syncthreads();
    } else if (SMEM) {
        __syncthreads(); // barrier init visible to the waiters
        mbar_wait(&bar, 0);
    }

    while (r < a.n_rows) {
        acc_t<NT> acc;
#pragma unroll
        for (int i = 0; i < NT; i++) acc.v[i] = 0.0f;
        if (VEC == 8) {
            uint32_t nz = lo + lane * 8;
            while (have_next) {
                uint32_t c[8];
#pragma unroll
                for (int i = 0; i < 8; i++) c[i] = c_next[i];
                val8<ValT> v = v_next;
                nz += 256;
                have_next = nz < hi;
                if (have_next) {
                    load_cols8<ColT>(col + nz, c_next);
                    v_next.load(val + nz);
                }
                step8<NT, SWZ>(acc, c, v, panel);
            }
        } else {
            for (uint32_t nz = lo + lane; nz < hi; nz += 32) nz1<NT, SWZ>(acc, val[nz], (uint32_t)col[nz], panel);
        }
        // next row's offsets and first loads before the reduction of this one
        uint32_t r2 = r + n_warps, lo2 = 0, hi2 = 0;
        if (r2 < a.n_rows) {
            lo2 = a.rowptr[r2], hi2 = a.rowptr[r2 + 1];
            if (VEC == 8) {
                uint32_t nz = lo2 + lane * 8;
                if (nz < hi2) {
                    load_cols8<ColT>(col + nz, c_next);
                    v_next.load(val + nz);
                    have_next = true;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < NT; i++) {
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) acc.v[i] += __shfl_xor_sync(0xffffffffu, acc.v[i], d);
        }
        if (lane == 0) {
            uint32_t row = a.rowid[r];
            row = (a.perm ? a.perm[row] : row) + a.begin_row;
            store_out<NT>((ValT *)a.C, row, acc);
        }
        r = r2, lo = lo2, hi = hi2;
    }
}

// ---- pipelined variant ---------------------------------------------------------------------------------------------
// v2 measured long_scoreboard as the dominant stall: one chunk of A per warp in flight is not enough to cover HBM
// latency (profiles/r1_row_v2_summary.md).  Here the warp's whole work -- its rows, chunk after chunk -- is one flat
// sequence, and D chunks (D x 1 KiB per warp) are kept in flight in registers across row boundaries.
template <typename ColT, typename ValT> struct chunk_t {
    uint4 c0, c1, v0, v1; // c1: u32 ids only; v1: f32 values only
    uint32_t row, flags;  // flags: 1 lane active, 2 last chunk of its row, 4 chunk exists
};

template <typename ColT> __device__ __forceinline__ void unpack_cols(const uint4 &c0, const uint4 &c1, uint32_t (&c)[8]);
template <> __device__ __forceinline__ void unpack_cols<uint16_t>(const uint4 &r, const uint4 &, uint32_t (&c)[8])
{
    c[0] = r.x & 0xffffu, c[1] = r.x >> 16, c[2] = r.y & 0xffffu, c[3] = r.y >> 16;
    c[4] = r.z & 0xffffu, c[5] = r.z >> 16, c[6] = r.w & 0xffffu, c[7] = r.w >> 16;
}
template <> __device__ __forceinline__ void unpack_cols<uint32_t>(const uint4 &r0, const uint4 &r1, uint32_t (&c)[8])
{
    c[0] = r0.x, c[1] = r0.y, c[2] = r0.z, c[3] = r0.w, c[4] = r1.x, c[5] = r1.y, c[6] = r1.z, c[7] = r1.w;
}

struct cursor_t {
    uint32_t r, hi, nz;   // row being loaded, its end, next chunk base
    uint32_t nlo, nhi;    // offsets of the row after it (prefetched)
};

template <int NT, typename ColT, typename ValT, int D> __global__ void __launch_bounds__(832, 1) k_b200_row_pipe(row_args a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    constexpr int ROWB = NT * (int)sizeof(ValT);
    constexpr bool SWZ = ROWB > 16 && ROWB <= 128;
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val;
    const int lane = threadIdx.x & 31;
    const uint32_t W = gridDim.x * (blockDim.x >> 5);

    if (!SWZ && threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_expect_tx(&bar, a.b_bytes);
        for (uint32_t off = 0; off < a.b_bytes; off += 32768u) {
            uint32_t n = a.b_bytes - off < 32768u ? a.b_bytes - off : 32768u;
            bulk_g2s(smem + off, (const unsigned char *)a.B + off, n, &bar);
        }
    }

    cursor_t cur;
    cur.r = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    cur.nz = 0, cur.hi = 0, cur.nlo = 0, cur.nhi = 0;
    if (cur.r < a.n_rows) cur.nz = a.rowptr[cur.r], cur.hi = a.rowptr[cur.r + 1];
    if (cur.r + W < a.n_rows) cur.nlo = a.rowptr[cur.r + W], cur.nhi = a.rowptr[cur.r + W + 1];

    auto load = [&](chunk_t<ColT, ValT> &st) {
        if (cur.r >= a.n_rows) {
            st.flags = 0;
            return;
        }
        uint32_t base = cur.nz + lane * 8;
        bool act = base < cur.hi;
        if (act) {
            st.c0 = ld_stream_v4(col + base);
            if (sizeof(ColT) == 4) st.c1 = ld_stream_v4(col + base + 4);
            st.v0 = ld_stream_v4(val + base);
            if (sizeof(ValT) == 4) st.v1 = ld_stream_v4(val + base + 4);
        }
        bool last = cur.nz + 256 >= cur.hi;
        st.row = cur.r;
        st.flags = 4u | (last ? 2u : 0u) | (act ? 1u : 0u);
        if (last) {
            cur.r += W;
            cur.nz = cur.nlo, cur.hi = cur.nhi;
            if (cur.r + W < a.n_rows) cur.nlo = a.rowptr[cur.r + W], cur.nhi = a.rowptr[cur.r + W + 1];
        } else {
            cur.nz += 256;
        }
    };

    chunk_t<ColT, ValT> st[D];
#pragma unroll
    for (int s = 0; s < D; s++) load(st[s]); // D chunks of A in flight before the panel has landed

    if (SWZ) {
        constexpr int P = ROWB / 16;
        const uint4 *src = (const uint4 *)a.B;
        uint32_t n_pieces = a.b_bytes >> 4;
        for (uint32_t x = threadIdx.x; x < n_pieces; x += blockDim.x) *(uint4 *)(smem + piece_offset<ROWB, true>(x / P, x % P)) = __ldg(src + x);
        __syncthreads();
    } else {
        __syncthreads();
        mbar_wait(&bar, 0);
    }

    acc_t<NT> acc;
#pragma unroll
    for (int i = 0; i < NT; i++) acc.v[i] = 0.0f;
    bool more = true;
    while (more) {
#pragma unroll
        for (int s = 0; s < D; s++) {
            const uint32_t fl = st[s].flags;
            if (!(fl & 4u)) {
                more = false;
                break;
            }
            if (fl & 1u) {
                uint32_t c[8];
                unpack_cols<ColT>(st[s].c0, st[s].c1, c);
                val8<ValT> v;
                if (sizeof(ValT) == 4) {
                    val8<float> *vf = (val8<float> *)&v;
                    vf->r0 = st[s].v0, vf->r1 = st[s].v1;
                } else {
                    ((val8<__half> *)&v)->r = st[s].v0;
                }
                step8<NT, SWZ>(acc, c, v, smem);
            }
            const uint32_t row_done = st[s].row;
            load(st[s]);
            if (fl & 2u) {
#pragma unroll
                for (int i = 0; i < NT; i++) {
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) acc.v[i] += __shfl_xor_sync(0xffffffffu, acc.v[i], d);
                }
                if (lane == 0) {
                    uint32_t row = a.rowid[row_done];
                    row = (a.perm ? a.perm[row] : row) + a.begin_row;
                    store_out<NT>((ValT *)a.C, row, acc);
                }
#pragma unroll
                for (int i = 0; i < NT; i++) acc.v[i] = 0.0f;
            }
        }
    }
}

// ---- B200 layout pass -------------------------------------------------------------------------------------------------
// Re-order the non-zeros INSIDE each row (SpMM is order-free inside a row) so that the 8 lanes of a quarter warp,
// which execute one 128-bit LDS together, hold column ids that differ mod 8.  Lane l of a warp step owns positions
// 8l..8l+7 of a 256-wide chunk; gather instruction j touches position 8l+j, so the 8 positions {64q+8t+j, t=0..7}
// form one conflict group.  Elements are bucketed by col % 8 and dealt round-robin over the groups.
// One thread per row; a one-time format pass (like the reference's interlance_storage step), not on the timed path.
template <typename ColT, typename ValT>
__global__ void k_rainbow_rows(const ColT *col, const ValT *val, ColT *col2, ValT *val2, ColT *tc, ValT *tv, const uint32_t *rowptr, uint32_t n_rows)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    uint32_t lo = rowptr[r], hi = rowptr[r + 1], L = hi - lo;
    if (L == 0) return;
    uint32_t Lf = (L / 256) * 256; // full 256-chunks, then a tail of m = (L-Lf)/8 lanes
    for (int part = 0; part < 2; part++) {
        uint32_t b = lo + (part ? Lf : 0), n = part ? L - Lf : Lf;
        if (n == 0) continue;
        uint32_t cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0}, start[8];
        for (uint32_t i = 0; i < n; i++) cnt[col[b + i] & 7]++;
        uint32_t s = 0;
        for (int k = 0; k < 8; k++) start[k] = s, s += cnt[k];
        for (uint32_t i = 0; i < n; i++) { // bucket-sorted copy
            uint32_t k = start[col[b + i] & 7]++;
            tc[b + k] = col[b + i], tv[b + k] = val[b + i];
        }
        if (part == 0) {
            uint32_t G = n / 8; // groups, all of size 8
            for (uint32_t k = 0; k < n; k++) {
                uint32_t g = k % G, t = k / G;
                uint32_t p = (g / 32) * 256 + ((g % 32) / 8) * 64 + t * 8 + (g % 8);
                col2[b + p] = tc[b + k], val2[b + p] = tv[b + k];
            }
        } else {
            uint32_t m = n / 8, k = 0; // active lanes in the tail chunk; quarter q has min(8, m-8q) of them
            for (uint32_t t = 0; t < 8; t++)
                for (uint32_t q = 0; q < 4; q++) {
                    if (8 * q + t >= m) continue;
                    for (uint32_t j = 0; j < 8; j++, k++) {
                        uint32_t p = q * 64 + t * 8 + j;
                        col2[b + p] = tc[b + k], val2[b + p] = tv[b + k];
                    }
                }
        }
    }
}


// ---- short rows: a GROUP of G lanes per row instead of a warp -------------------------------------------------------------------------
// One warp per row (k_b200_row) leaves most lanes idle on matrices like BASELINE's C1 (IG5-18: 37 non-zeros per row) and walks its rows
// one after the other: 12 rows per warp, each a chain of dependent loads (row pointer -> ids / values -> rows of B -> shuffles -> store),
// 26.7 us for a 14 MB problem.  Here 32 / G rows are in flight per warp: lane j of a group takes the row's non-zeros j, j + G, ... with all
// NT dense columns in registers (a row of B is one or two 128-bit loads), a halving butterfly leaves lane j with columns
// j * NT / G ... of the row sum (NT - NT / G shuffles instead of NT * log2 G), and the group stores the row of C as one coalesced piece.
// Groups take rows gid, gid + groups, ...: after sort_operator (rows by length, descending) the rows of a warp are equally long.
// The dense matrix is read through L1 / L2 (this variant is for panels that do not fit shared memory).
template <int NT, typename ColT, typename ValT, int G> __global__ void __launch_bounds__(256) k_b200_row_short(row_args a)
{
    static_assert(NT % G == 0 || G % NT == 0, "columns per lane");
    constexpr int PER = NT >= G ? NT / G : 1; // columns a lane owns after the reduction (G > NT: only the first NT lanes own one)
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val, *B = (const ValT *)a.B;
    ValT *C = (ValT *)a.C;
    const int lane = threadIdx.x & 31, j = lane % G;
    const unsigned mask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane & ~(G - 1)));
    const uint32_t groups = gridDim.x * (256 / G), gid = (blockIdx.x * 256 + threadIdx.x) / G;
    constexpr int VPL = 16 / sizeof(ValT); // values per 128-bit load
    for (uint32_t r = gid; r < a.n_rows; r += groups) {
        const uint32_t lo = a.rowptr[r], hi = a.rowptr[r + 1];
        float acc[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) acc[i] = 0.0f;
        for (uint32_t k = lo + j; k < hi; k += 2 * G) { // two non-zeros per lane and round: their loads go out together
            const bool two = k + G < hi;
            const uint32_t c0 = (uint32_t)col[k], c1 = two ? (uint32_t)col[k + G] : c0;
            const float v0 = to_f32(val[k]), v1 = two ? to_f32(val[k + G]) : 0.0f;
            uint4 b0[NT / VPL], b1[NT / VPL];
#pragma unroll
            for (int q = 0; q < NT / VPL; q++) b0[q] = __ldg((const uint4 *)(B + (size_t)c0 * NT) + q), b1[q] = __ldg((const uint4 *)(B + (size_t)c1 * NT) + q);
#pragma unroll
            for (int q = 0; q < NT / VPL; q++) {
                const ValT *x0 = (const ValT *)&b0[q], *x1 = (const ValT *)&b1[q];
#pragma unroll
                for (int t = 0; t < VPL; t++) acc[q * VPL + t] = fmaf(v1, to_f32(x1[t]), fmaf(v0, to_f32(x0[t]), acc[q * VPL + t]));
            }
        }
        // halving butterfly: at distance d the lane keeps the half of its columns its bit d selects and adds the partner's
        int n = NT;
#pragma unroll
        for (int d = G / 2; d >= 1; d >>= 1) {
            if (n > 1) {
                n >>= 1;
                const bool upper = (j & d) != 0;
#pragma unroll
                for (int i = 0; i < NT / 2; i++) {
                    if (i < n) {
                        const float keep = upper ? acc[i + n] : acc[i], give = upper ? acc[i] : acc[i + n];
                        acc[i] = keep + __shfl_xor_sync(mask, give, d, G);
                    }
                }
            } else { // more lanes than columns left: plain sum, the lanes of one column end up with the same total
                acc[0] += __shfl_xor_sync(mask, acc[0], d, G);
            }
        }
        const uint32_t row = a.rowid[r];
        ValT *dst = C + (size_t)((a.perm ? a.perm[row] : row) + a.begin_row) * NT;
        if (NT >= G) {
#pragma unroll
            for (int i = 0; i < PER; i++) dst[j * PER + i] = from_f32<ValT>(acc[i]);
        } else { // G > NT: after log2(NT) halving steps lane j holds column (j / (G / NT)); one lane per column stores
            if (j % (G / NT) == 0) dst[j / (G / NT)] = from_f32<ValT>(acc[0]);
        }
    }
}

} // namespace row
} // namespace gsd
