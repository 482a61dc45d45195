// B200-native row kernels (GS_FAMILY_B200_ROW_SMEM): the throughput path for the pruned-weight
// configs.  They consume the same arrays a row-direction chain of the reference produces
// (nz_col_indices, nz_vals, first_row_indices / first_nz_indices of the BMT, BMW or BMTB level with
// row block 1, optional original_nz_row_indices) -- i.e. what the total_BMT/total_warp/total_block
// tokens read (total_BMT_result_reduce_to_one_register_token.cc:828-892) -- but map the work for
// sm_100a:
//   * persistent CTAs, one per SM; one warp per row; the number of warps per CTA is chosen so that
//     rows / warps is (nearly) an integer -- at 11008 rows, 25 warps x 148 SMs do exactly 3 rows each
//     instead of 2.3 with a full 32-warp CTA (rows are length-sorted by sort_operator, so a strided
//     assignment balances itself);
//   * each lane owns 8 consecutive non-zeros per step: one 128-bit ld.global.nc of 8 u16 column
//     ids + one of 8 f16 values (the reference's Load() of a sparse-coarsen vector,
//     cuda_code/kernel_lib.hpp:1006-1148), next step's loads issued before this step's math;
//   * the dense B panel (K x N) is staged once per CTA in shared memory -- a bulk async copy
//     (cp.async.bulk -> UBLKCP, completion on an mbarrier) when a row is one 16-byte piece, a swizzled
//     cooperative copy otherwise -- while the first A loads are already in flight; gathers are 128-bit LDS;
//   * B200 layout pass: the non-zeros inside each row are re-ordered once so that the 8 lanes of a
//     quarter warp gather from 8 distinct bank groups (v1 measured 61 % of the LDS wavefronts as bank
//     conflicts, profiles/r1_row_v1_summary.md);
//   * fp16 x fp16 products accumulate in fp32 with the mixed-precision FMA (fma.rn.f32.f16 -> FHFMA);
//   * rows are disjoint, so C is written with plain 128-bit stores (no atomics, no memset).
// HBM-bound byte work: nothing here is shaped like a GEMM.
#include "b200_row.cuh"

#include <cstdio>
#include <cstdlib>
#include <new>

namespace gsd {

struct b200_row_plan {
    const uint32_t *rowptr = nullptr; // first_nz_indices of the row level (n_rows + 1 entries)
    const uint32_t *rowid = nullptr;  // first_row_indices of the same level
    uint32_t n_rows = 0;
    int vec = 1; // 8: rows padded to 8 (sparse coarsen 8), vector loads; 1: scalar
    bool smem_b = false;
    int grid = 0, threads = 512;
    size_t smem_bytes = 0;
    const void *fn = nullptr;
    char name[96] = {0};
    void *col2 = nullptr, *val2 = nullptr; // B200 layout (owned); the reference-order arrays stay untouched
};

namespace {

using namespace row;

template <int NT, typename ColT, typename ValT> const void *pick_fn(int vec, bool smem)
{
    if (vec == 8) return smem ? (const void *)k_b200_row<NT, ColT, ValT, 8, true> : (const void *)k_b200_row<NT, ColT, ValT, 8, false>;
    return smem ? (const void *)k_b200_row<NT, ColT, ValT, 1, true> : (const void *)k_b200_row<NT, ColT, ValT, 1, false>;
}

template <int NT, typename ColT, typename ValT> const void *pick_pipe(int depth)
{
    switch (depth) {
    case 2: return (const void *)k_b200_row_pipe<NT, ColT, ValT, 2>;
    case 3: return (const void *)k_b200_row_pipe<NT, ColT, ValT, 3>;
    case 4: return (const void *)k_b200_row_pipe<NT, ColT, ValT, 4>;
    case 6: return (const void *)k_b200_row_pipe<NT, ColT, ValT, 6>;
    default: return nullptr;
    }
}

template <typename ColT, typename ValT> const void *pick_pipe_nt(int N, int depth)
{
    switch (N) {
    case 8: return pick_pipe<8, ColT, ValT>(depth);
    case 16: return pick_pipe<16, ColT, ValT>(depth);
    default: return nullptr; // wider panels do not fit shared memory in fp16 x 4096 rows; the generic kernel handles them
    }
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

template <typename ColT, typename ValT> cudaError_t rainbow(const launch_params &p, b200_row_plan *pl, cudaStream_t s)
{
    size_t cb = p.a.nnz * sizeof(ColT) + 16, vb = p.a.nnz * sizeof(ValT) + 16;
    void *tc = nullptr, *tv = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc(&pl->col2, cb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&pl->val2, vb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&tc, cb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&tv, vb)) != cudaSuccess) {
        cudaFree(tc);
        return e;
    }
    // start from a plain copy so that anything the pass does not touch (nothing, by construction) stays valid
    cudaMemcpyAsync(pl->col2, p.a.col, cb - 16, cudaMemcpyDeviceToDevice, s);
    cudaMemcpyAsync(pl->val2, p.a.val, vb - 16, cudaMemcpyDeviceToDevice, s);
    k_rainbow_rows<ColT, ValT><<<(pl->n_rows + 63) / 64, 64, 0, s>>>((const ColT *)p.a.col, (const ValT *)p.a.val, (ColT *)pl->col2, (ValT *)pl->val2,
                                                                      (ColT *)tc, (ValT *)tv, pl->rowptr, pl->n_rows);
    e = cudaStreamSynchronize(s);
    cudaFree(tc);
    cudaFree(tv);
    return e != cudaSuccess ? e : cudaGetLastError();
}

int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

} // namespace

cudaError_t b200_row_prepare(const launch_params &p, int sm_count, b200_row_plan **out, cudaStream_t s)
{
    *out = nullptr;
    b200_row_plan *pl = new (std::nothrow) b200_row_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    auto fail = [&](cudaError_t e) {
        b200_row_free(pl);
        return e;
    };
    // the row level: BMT, BMW or BMTB arrays with one row per block
    if (p.a.t_first_nz && p.a.t_first_row && !p.a.t_nz_relative) pl->rowptr = p.a.t_first_nz, pl->rowid = p.a.t_first_row, pl->n_rows = p.a.n_t;
    else if (p.a.w_first_nz && p.a.w_first_row) pl->rowptr = p.a.w_first_nz, pl->rowid = p.a.w_first_row, pl->n_rows = p.a.n_w;
    else if (p.a.tb_first_nz && p.a.tb_first_row) pl->rowptr = p.a.tb_first_nz, pl->rowid = p.a.tb_first_row, pl->n_rows = p.a.n_tb;
    else return fail(cudaErrorInvalidValue);
    size_t esz = p.a.val_is_f16 ? 2 : 4;
    size_t panel = (size_t)p.a.cols * p.N * esz;
    pl->smem_b = panel <= 200 * 1024 && panel % 16 == 0 && !env_int("GS_ROW_NO_SMEM", 0);
    pl->smem_bytes = pl->smem_b ? panel : 0;
    pl->vec = p.vector_width; // the engine passes the row alignment here: 8 when every row offset is a multiple of 8
    const void *fn = nullptr;
    if (p.a.col_is_u16) fn = p.a.val_is_f16 ? pick_nt<uint16_t, __half>(p.N, pl->vec, pl->smem_b) : pick_nt<uint16_t, float>(p.N, pl->vec, pl->smem_b);
    else fn = p.a.val_is_f16 ? pick_nt<uint32_t, __half>(p.N, pl->vec, pl->smem_b) : pick_nt<uint32_t, float>(p.N, pl->vec, pl->smem_b);
    if (!fn) return fail(cudaErrorInvalidValue); // dense width not instantiated
    // pipelined variant: rows padded to 8, panel in shared memory, D chunks of A per warp in flight
    // (opt-in: measured slower than the plain kernel at every depth on C2, profiles/r1_row_v2_summary.md)
    int depth = env_int("GS_ROW_DEPTH", 1), max_w = 32;
    if (pl->vec == 8 && pl->smem_b && depth > 1) {
        const void *pf = nullptr;
        if (p.a.col_is_u16) pf = p.a.val_is_f16 ? pick_pipe_nt<uint16_t, __half>(p.N, depth) : pick_pipe_nt<uint16_t, float>(p.N, depth);
        else pf = p.a.val_is_f16 ? pick_pipe_nt<uint32_t, __half>(p.N, depth) : pick_pipe_nt<uint32_t, float>(p.N, depth);
        if (pf) fn = pf, max_w = 26; // __launch_bounds__(832): room for the in-flight chunks in registers
        else depth = 1;
    } else depth = 1;
    pl->fn = fn;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_bytes);
    if (e != cudaSuccess) return fail(e);

    // warps per CTA: the candidate in [16, 32] that wastes the fewest row slots; ties go to more warps
    int best_w = max_w;
    double best_eff = 0;
    for (int w = max_w; w >= 16; w--) {
        uint64_t warps = (uint64_t)sm_count * w;
        uint64_t rounds = (pl->n_rows + warps - 1) / warps;
        double eff = (double)pl->n_rows / (double)(rounds * warps);
        if (eff > best_eff + 1e-9) best_eff = eff, best_w = w;
    }
    best_w = env_int("GS_ROW_WARPS", best_w);
    pl->threads = best_w * 32;
    pl->grid = sm_count;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, pl->threads, pl->smem_bytes);
    if (e != cudaSuccess || occ < 1) return fail(e != cudaSuccess ? e : cudaErrorInvalidConfiguration);

    bool do_rainbow = pl->vec == 8 && pl->smem_b && !env_int("GS_ROW_NO_RAINBOW", 0);
    if (do_rainbow) {
        if (p.a.col_is_u16) e = p.a.val_is_f16 ? rainbow<uint16_t, __half>(p, pl, s) : rainbow<uint16_t, float>(p, pl, s);
        else e = p.a.val_is_f16 ? rainbow<uint32_t, __half>(p, pl, s) : rainbow<uint32_t, float>(p, pl, s);
        if (e != cudaSuccess) return fail(e);
    }
    snprintf(pl->name, sizeof pl->name, "k_b200_row%s<N%d,%s,%s,vec%d,%s%s,%dw,d%d>", depth > 1 ? "_pipe" : "", p.N, p.a.col_is_u16 ? "u16" : "u32", p.a.val_is_f16 ? "f16" : "f32", pl->vec,
             pl->smem_b ? "smem" : "gmem", do_rainbow ? "+rainbow" : "", best_w, depth);
    *out = pl;
    return cudaSuccess;
}

cudaError_t b200_row_launch(const b200_row_plan *pl, const launch_params &p, cudaStream_t s)
{
    row::row_args a;
    a.col = pl->col2 ? pl->col2 : p.a.col, a.val = pl->val2 ? pl->val2 : p.a.val, a.B = p.B, a.C = p.C;
    a.rowptr = pl->rowptr, a.rowid = pl->rowid, a.perm = p.a.perm;
    a.n_rows = pl->n_rows, a.begin_row = p.a.begin_row, a.K = p.a.cols;
    a.b_bytes = (uint32_t)pl->smem_bytes;
    void *args[] = {&a};
    return cudaLaunchKernel(pl->fn, dim3(pl->grid), dim3(pl->threads), args, pl->smem_bytes, s);
}

void b200_row_free(b200_row_plan *pl)
{
    if (!pl) return;
    if (pl->col2) cudaFree(pl->col2);
    if (pl->val2) cudaFree(pl->val2);
    delete pl;
}
const char *b200_row_kernel_name(const b200_row_plan *pl) { return pl ? pl->name : ""; }

} // namespace gsd
