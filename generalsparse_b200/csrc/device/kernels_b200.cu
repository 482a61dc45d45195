// B200-native row kernels (GS_FAMILY_B200_ROW_SMEM): the throughput path for the pruned-weight
// configs.  They consume the same arrays a row-direction chain of the reference produces
// (nz_col_indices, nz_vals, first_row_indices / first_nz_indices of the BMT, BMW or BMTB level with
// row block 1, optional original_nz_row_indices) -- i.e. what total_BMT/total_warp/total_block
// tokens read (total_BMT_result_reduce_to_one_register_token.cc:828-892) -- but map the work for
// sm_100a:
//   * persistent CTAs (grid = SMs x resident CTAs), one warp per row, rows taken grid-stride
//     (after sort_operator the rows are length-sorted, so the stride balances itself);
//   * each lane owns 8 consecutive non-zeros per step: one 128-bit ld.global.nc of 8 u16 column
//     ids + one of 8 f16 values (the reference's Load() of a sparse-coarsen vector,
//     cuda_code/kernel_lib.hpp:1006-1148), next step's loads issued before this step's math;
//   * the dense B panel (K x N) is staged once per CTA in shared memory by a bulk async copy
//     (cp.async.bulk, completion on an mbarrier) while the first A loads are already in flight;
//     gathers are then 128-bit LDS;
//   * fp16 x fp16 products accumulate in fp32 with the mixed-precision FMA (fma.rn.f32.f16 -> FHFMA);
//   * rows are disjoint, so C is written with plain 128-bit stores (no atomics, no memset).
// HBM-bound byte work: nothing here is shaped like a GEMM; the tensor-core path is separate.
#include "common.cuh"

#include <new>

namespace gsd {

struct b200_row_plan {
    const uint32_t *rowptr = nullptr;  // first_nz_indices of the row level (n_rows + 1 entries)
    const uint32_t *rowid = nullptr;   // first_row_indices of the same level
    uint32_t n_rows = 0;
    int vec = 1;                       // 8: rows padded to 8 (sparse coarsen 8), vector loads; 1: scalar
    bool smem_b = false;
    int grid = 0, threads = 512;
    size_t smem_bytes = 0;
    const void *fn = nullptr;
    const char *name = "";
};

namespace {

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
__device__ __forceinline__ void fma_h(float &acc, unsigned short a, unsigned short b)
{
    asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(acc) : "h"(a), "h"(b));
}
__device__ __forceinline__ void split_h2(uint32_t x, unsigned short &lo, unsigned short &hi)
{
    asm("mov.b32 {%0,%1}, %2;" : "=h"(lo), "=h"(hi) : "r"(x));
}

template <int NT> struct acc_t {
    float v[NT];
};

// one non-zero (value bits `a`, dense row at `brow`) into NT accumulators -- half storage
template <int NT> __device__ __forceinline__ void nz_fma_h(acc_t<NT> &acc, unsigned short a, const uint4 *brow)
{
#pragma unroll
    for (int q = 0; q < NT / 8; q++) {
        uint4 b = brow[q];
        unsigned short h0, h1;
        split_h2(b.x, h0, h1), fma_h(acc.v[q * 8 + 0], a, h0), fma_h(acc.v[q * 8 + 1], a, h1);
        split_h2(b.y, h0, h1), fma_h(acc.v[q * 8 + 2], a, h0), fma_h(acc.v[q * 8 + 3], a, h1);
        split_h2(b.z, h0, h1), fma_h(acc.v[q * 8 + 4], a, h0), fma_h(acc.v[q * 8 + 5], a, h1);
        split_h2(b.w, h0, h1), fma_h(acc.v[q * 8 + 6], a, h0), fma_h(acc.v[q * 8 + 7], a, h1);
    }
}
// float storage
template <int NT> __device__ __forceinline__ void nz_fma_f(acc_t<NT> &acc, float a, const float4 *brow)
{
#pragma unroll
    for (int q = 0; q < NT / 4; q++) {
        float4 b = brow[q];
        acc.v[q * 4 + 0] = fmaf(a, b.x, acc.v[q * 4 + 0]);
        acc.v[q * 4 + 1] = fmaf(a, b.y, acc.v[q * 4 + 1]);
        acc.v[q * 4 + 2] = fmaf(a, b.z, acc.v[q * 4 + 2]);
        acc.v[q * 4 + 3] = fmaf(a, b.w, acc.v[q * 4 + 3]);
    }
}

template <typename ColT> struct col8 { // 8 column ids of one lane step
    uint32_t c[8];
};
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

template <int NT> __device__ __forceinline__ void step8(acc_t<NT> &acc, const uint32_t (&c)[8], const val8<__half> &v, const __half *Bp)
{
    const uint32_t w[4] = {v.r.x, v.r.y, v.r.z, v.r.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
        unsigned short a0, a1;
        split_h2(w[j], a0, a1);
        nz_fma_h<NT>(acc, a0, (const uint4 *)(Bp + (size_t)c[2 * j] * NT));
        nz_fma_h<NT>(acc, a1, (const uint4 *)(Bp + (size_t)c[2 * j + 1] * NT));
    }
}
template <int NT> __device__ __forceinline__ void step8(acc_t<NT> &acc, const uint32_t (&c)[8], const val8<float> &v, const float *Bp)
{
    const uint32_t w[8] = {v.r0.x, v.r0.y, v.r0.z, v.r0.w, v.r1.x, v.r1.y, v.r1.z, v.r1.w};
#pragma unroll
    for (int j = 0; j < 8; j++) nz_fma_f<NT>(acc, __uint_as_float(w[j]), (const float4 *)(Bp + (size_t)c[j] * NT));
}

template <int NT> __device__ __forceinline__ void nz1(acc_t<NT> &acc, __half v, uint32_t c, const __half *Bp)
{
    nz_fma_h<NT>(acc, __half_as_ushort(v), (const uint4 *)(Bp + (size_t)c * NT));
}
template <int NT> __device__ __forceinline__ void nz1(acc_t<NT> &acc, float v, uint32_t c, const float *Bp)
{
    nz_fma_f<NT>(acc, v, (const float4 *)(Bp + (size_t)c * NT));
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

// NT = N (dense width), VEC = 8 (rows padded to multiples of 8) or 1, SMEM = panel staged in shared memory
template <int NT, typename ColT, typename ValT, int VEC, bool SMEM> __global__ void __launch_bounds__(512) k_b200_row(row_args a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    const ColT *col = (const ColT *)a.col;
    const ValT *val = (const ValT *)a.val;
    const ValT *Bp = (const ValT *)a.B;
    const int lane = threadIdx.x & 31;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const uint32_t warp0 = blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
    const uint32_t n_warps = gridDim.x * warps_per_cta;

    if (SMEM) {
        if (threadIdx.x == 0) {
            mbar_init(&bar, 1);
            mbar_expect_tx(&bar, a.b_bytes);
            // one bulk copy per 32 KiB piece, all completing on the same barrier
            for (uint32_t off = 0; off < a.b_bytes; off += 32768u) {
                uint32_t n = a.b_bytes - off < 32768u ? a.b_bytes - off : 32768u;
                bulk_g2s(smem + off, (const unsigned char *)a.B + off, n, &bar);
            }
        }
        Bp = (const ValT *)smem;
    }

    uint32_t r = warp0;
    uint32_t lo = 0, hi = 0;
    if (r < a.n_rows) lo = a.rowptr[r], hi = a.rowptr[r + 1];
    // first A loads go out before the panel has landed
    uint32_t c_next[8];
    val8<ValT> v_next;
    bool have_next = false;
    if (VEC == 8) {
        uint32_t nz = lo + lane * 8;
        if (nz < hi) {
            load_cols8<ColT>(col + nz, c_next);
            v_next.load(val + nz);
            have_next = true;
        }
    }
    if (SMEM) {
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
                step8<NT>(acc, c, v, Bp);
            }
        } else {
            for (uint32_t nz = lo + lane; nz < hi; nz += 32) nz1<NT>(acc, val[nz], (uint32_t)col[nz], Bp);
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

template <int NT, typename ColT, typename ValT> const void *pick_fn(int vec, bool smem)
{
    if (vec == 8) return smem ? (const void *)k_b200_row<NT, ColT, ValT, 8, true> : (const void *)k_b200_row<NT, ColT, ValT, 8, false>;
    return smem ? (const void *)k_b200_row<NT, ColT, ValT, 1, true> : (const void *)k_b200_row<NT, ColT, ValT, 1, false>;
}

template <typename ColT, typename ValT> const void *pick_nt(int N, int vec, bool smem)
{
    switch (N) {
    case 8: return pick_fn<8, ColT, ValT>(vec, smem);
    case 16: return pick_fn<16, ColT, ValT>(vec, smem);
    case 32: return pick_fn<32, ColT, ValT>(vec, smem);
    default: return nullptr;
    }
}

} // namespace

cudaError_t b200_row_prepare(const launch_params &p, int sm_count, b200_row_plan **out, cudaStream_t)
{
    *out = nullptr;
    b200_row_plan *pl = new (std::nothrow) b200_row_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    // the row level: BMT, BMW or BMTB arrays with one row per block
    if (p.a.t_first_nz && p.a.t_first_row && !p.a.t_nz_relative) pl->rowptr = p.a.t_first_nz, pl->rowid = p.a.t_first_row, pl->n_rows = p.a.n_t;
    else if (p.a.w_first_nz && p.a.w_first_row) pl->rowptr = p.a.w_first_nz, pl->rowid = p.a.w_first_row, pl->n_rows = p.a.n_w;
    else if (p.a.tb_first_nz && p.a.tb_first_row) pl->rowptr = p.a.tb_first_nz, pl->rowid = p.a.tb_first_row, pl->n_rows = p.a.n_tb;
    else {
        delete pl;
        return cudaErrorInvalidValue;
    }
    size_t esz = p.a.val_is_f16 ? 2 : 4;
    size_t panel = (size_t)p.a.cols * p.N * esz;
    pl->smem_b = panel <= 200 * 1024 && panel % 16 == 0;
    pl->smem_bytes = pl->smem_b ? panel : 0;
    pl->vec = p.vector_width; // the engine passes the row alignment here: 8 when every row offset is a multiple of 8
    const void *fn = nullptr;
    if (p.a.col_is_u16) fn = p.a.val_is_f16 ? pick_nt<uint16_t, __half>(p.N, pl->vec, pl->smem_b) : pick_nt<uint16_t, float>(p.N, pl->vec, pl->smem_b);
    else fn = p.a.val_is_f16 ? pick_nt<uint32_t, __half>(p.N, pl->vec, pl->smem_b) : pick_nt<uint32_t, float>(p.N, pl->vec, pl->smem_b);
    if (!fn) {
        delete pl;
        return cudaErrorInvalidValue; // dense width not instantiated
    }
    pl->fn = fn;
    pl->name = pl->smem_b ? (pl->vec == 8 ? "k_b200_row<vec8,smem>" : "k_b200_row<vec1,smem>") : (pl->vec == 8 ? "k_b200_row<vec8,gmem>" : "k_b200_row<vec1,gmem>");
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_bytes);
    if (e != cudaSuccess) {
        delete pl;
        return e;
    }
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, pl->threads, pl->smem_bytes);
    if (e != cudaSuccess || occ < 1) {
        delete pl;
        return e != cudaSuccess ? e : cudaErrorInvalidConfiguration;
    }
    pl->grid = sm_count * occ;
    *out = pl;
    return cudaSuccess;
}

cudaError_t b200_row_launch(const b200_row_plan *pl, const launch_params &p, cudaStream_t s)
{
    row_args a;
    a.col = p.a.col, a.val = p.a.val, a.B = p.B, a.C = p.C;
    a.rowptr = pl->rowptr, a.rowid = pl->rowid, a.perm = p.a.perm;
    a.n_rows = pl->n_rows, a.begin_row = p.a.begin_row, a.K = p.a.cols;
    a.b_bytes = (uint32_t)pl->smem_bytes;
    void *args[] = {&a};
    return cudaLaunchKernel(pl->fn, dim3(pl->grid), dim3(pl->threads), args, pl->smem_bytes, s);
}

void b200_row_free(b200_row_plan *pl) { delete pl; }
const char *b200_row_kernel_name(const b200_row_plan *pl) { return pl ? pl->name : ""; }

} // namespace gsd
