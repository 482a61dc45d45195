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
#include "b200_stream.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>
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
    // stream schedule (b200_stream.cuh): per-warp ranges of the non-zero stream, computed once at plan time
    bool stream = false;
    uint32_t *w_nz = nullptr, *w_row = nullptr; // device, grid x (warps + 1) each (owned)
    uint32_t bulk_panel = 0, pf = 0;
    size_t table_bytes = 0; // shared memory behind the panel: the largest CTA row table
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

template <int NT, typename ColT, typename ValT> const void *pick_short_g(int G)
{
    switch (G) {
    case 4: return (const void *)row::k_b200_row_short<NT, ColT, ValT, 4>;
    case 8: return (const void *)row::k_b200_row_short<NT, ColT, ValT, 8>;
    case 16: return (const void *)row::k_b200_row_short<NT, ColT, ValT, 16>;
    default: return nullptr;
    }
}
template <typename ColT, typename ValT> const void *pick_short(int N, int G)
{
    switch (N) {
    case 8: return pick_short_g<8, ColT, ValT>(G);
    case 16: return pick_short_g<16, ColT, ValT>(G);
    case 32: return pick_short_g<32, ColT, ValT>(G);
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

// ---- stream schedule: kernel pickers, work partition, layout pass ------------------------------------------------
template <int NT, typename ColT, typename ValT, int CTAS, bool PRE = false> const void *pick_stream_d(int depth)
{
    switch (depth) {
    case 1: return (const void *)k_b200_stream<NT, ColT, ValT, 1, CTAS, PRE>;
    case 2: return (const void *)k_b200_stream<NT, ColT, ValT, 2, CTAS, PRE>;
    default: return nullptr;
    }
}
template <typename ColT, typename ValT> const void *pick_stream(int N, int depth, int ctas)
{
    if (N == 8) return ctas == 1 ? pick_stream_d<8, ColT, ValT, 1>(depth) : pick_stream_d<8, ColT, ValT, 2>(depth);
    if (N == 16) return ctas == 1 ? pick_stream_d<16, ColT, ValT, 1>(depth) : pick_stream_d<16, ColT, ValT, 2>(depth);
    return nullptr;
}
// PRE variant (pre-scaled column ids): N = 8, u16 ids, fp16, K * 16 <= 65536
const void *pick_stream_pre(int depth, int ctas)
{
    return ctas == 1   ? pick_stream_d<8, uint16_t, __half, 1, true>(depth)
           : ctas == 2 ? pick_stream_d<8, uint16_t, __half, 2, true>(depth)
                       : pick_stream_d<8, uint16_t, __half, 3, true>(depth);
}

// CTA c owns the rows [cta_row[c], cta_row[c+1]): cut points at the row boundary nearest to c/n of the non-zeros.
// Inside a CTA the 256-wide steps are dealt in equal runs to the warps; w_row holds, for every warp boundary, the
// first row that starts at or after it (binary search over the CTA's rows).
void stream_partition(const std::vector<uint32_t> &rp, uint32_t n_rows, int n_cta, int W, std::vector<uint32_t> &cta_row, std::vector<uint32_t> &w_nz,
                      std::vector<uint32_t> &w_row)
{
    cta_row.assign(n_cta + 1, 0);
    const uint64_t first = rp[0], total = rp[n_rows] - first;
    for (int c = 1; c < n_cta; c++) {
        uint64_t target = first + total * (uint64_t)c / (uint64_t)n_cta;
        uint32_t r = (uint32_t)(std::lower_bound(rp.begin(), rp.begin() + n_rows + 1, (uint32_t)target) - rp.begin());
        if (r > 0 && target - rp[r - 1] < rp[r] - target) r--;
        cta_row[c] = std::max(std::min(r, n_rows), cta_row[c - 1]);
    }
    cta_row[n_cta] = n_rows;
    w_nz.assign((size_t)n_cta * (W + 1), 0);
    w_row.assign((size_t)n_cta * (W + 1), 0);
    for (int c = 0; c < n_cta; c++) {
        const uint32_t r0 = cta_row[c], r1 = cta_row[c + 1], nz0 = rp[r0], nz1 = rp[r1];
        const uint32_t steps = (nz1 - nz0 + 255) / 256, per = (steps + W - 1) / W;
        for (int w = 0; w <= W; w++) {
            uint64_t st = std::min<uint64_t>(steps, (uint64_t)w * per);
            uint32_t nz = (uint32_t)std::min<uint64_t>(nz1, nz0 + st * 256);
            if (w == W) nz = nz1;
            w_nz[(size_t)c * (W + 1) + w] = nz;
            uint32_t r = (uint32_t)(std::lower_bound(rp.begin() + r0, rp.begin() + r1, nz) - rp.begin());
            w_row[(size_t)c * (W + 1) + w] = w == W ? r1 : r;
        }
    }
}

template <typename ColT, typename ValT>
cudaError_t rainbow_stream(const launch_params &p, b200_row_plan *pl, const std::vector<uint32_t> &cta_row, uint32_t scale, cudaStream_t s)
{
    size_t cb = p.a.nnz * sizeof(ColT) + 16, vb = p.a.nnz * sizeof(ValT) + 16;
    void *tc = nullptr, *tv = nullptr;
    uint32_t *d_cta = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc(&pl->col2, cb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&pl->val2, vb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&tc, cb)) == cudaSuccess && (e = cudaMalloc(&tv, vb)) == cudaSuccess && (e = cudaMalloc(&d_cta, cta_row.size() * 4)) == cudaSuccess) {
        cudaMemcpyAsync(pl->col2, p.a.col, cb - 16, cudaMemcpyDeviceToDevice, s);
        cudaMemcpyAsync(pl->val2, p.a.val, vb - 16, cudaMemcpyDeviceToDevice, s);
        cudaMemcpyAsync(d_cta, cta_row.data(), cta_row.size() * 4, cudaMemcpyHostToDevice, s);
        k_rainbow_stream<ColT, ValT><<<(pl->n_rows + 63) / 64, 64, 0, s>>>((const ColT *)p.a.col, (const ValT *)p.a.val, (ColT *)pl->col2, (ValT *)pl->val2,
                                                                            (ColT *)tc, (ValT *)tv, pl->rowptr, pl->n_rows, d_cta, (uint32_t)cta_row.size() - 1, scale);
        e = cudaStreamSynchronize(s);
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    cudaFree(tc), cudaFree(tv), cudaFree(d_cta);
    return e;
}

// Returns cudaErrorNotSupported when the stream schedule does not apply (the caller falls back to one warp per row).
cudaError_t stream_prepare(const launch_params &p, int sm_count, b200_row_plan *pl, cudaStream_t s)
{
    const int depth = env_int("GS_STREAM_DEPTH", 1), shape = env_int("GS_STREAM_CTAS", 2);
    if (shape < 1 || shape > 3) return cudaErrorNotSupported; // (3 CTAs of 10 warps: pre-scaled variant only)
    const int ctas = shape, max_w = shape == 1 ? 32 : shape == 2 ? 16 : 10;
    const int W = env_int("GS_STREAM_WARPS", max_w);
    if (W < 1 || W > max_w) return cudaErrorNotSupported;
    const void *fn = nullptr;
    if (p.a.col_is_u16) fn = p.a.val_is_f16 ? pick_stream<uint16_t, __half>(p.N, depth, shape) : pick_stream<uint16_t, float>(p.N, depth, shape);
    else fn = p.a.val_is_f16 ? pick_stream<uint32_t, __half>(p.N, depth, shape) : pick_stream<uint32_t, float>(p.N, depth, shape);
    if (!fn) return cudaErrorNotSupported;
    if (pl->smem_bytes * ctas + 8192 > 227 * 1024) return cudaErrorNotSupported;

    std::vector<uint32_t> rp(pl->n_rows + 1), cta_row, w_nz, w_row;
    cudaError_t e = cudaMemcpyAsync(rp.data(), pl->rowptr, rp.size() * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return e;
    const int n_cta = sm_count * ctas;
    stream_partition(rp, pl->n_rows, n_cta, W, cta_row, w_nz, w_row);
    // the CTA's row table (rowptr slice + output rows) sits behind the panel in shared memory
    uint32_t max_rows = 0;
    for (int c = 0; c < n_cta; c++) max_rows = std::max(max_rows, cta_row[c + 1] - cta_row[c]);
    pl->table_bytes = ((size_t)(2 * max_rows + 1) * 4 + (size_t)max_rows * p.N * 4 + 15) / 16 * 16; // + the fp32 C tile
    if (pl->table_bytes > 16384 || (pl->smem_bytes + pl->table_bytes + 2048) * ctas > 227 * 1024) return cudaErrorNotSupported;
    const size_t dyn = pl->smem_bytes + pl->table_bytes;
    if ((e = cudaMalloc(&pl->w_nz, w_nz.size() * 4)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&pl->w_row, w_row.size() * 4)) != cudaSuccess) return e;
    cudaMemcpyAsync(pl->w_nz, w_nz.data(), w_nz.size() * 4, cudaMemcpyHostToDevice, s);
    cudaMemcpyAsync(pl->w_row, w_row.data(), w_row.size() * 4, cudaMemcpyHostToDevice, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return e;

    const bool do_rainbow = !env_int("GS_ROW_NO_RAINBOW", 0);
    // pre-scaled ids: the layout pass stores id * 16 (the byte offset of the dense row in the panel) when that fits u16
    const bool pre = do_rainbow && p.N == 8 && p.a.col_is_u16 && p.a.val_is_f16 && (uint64_t)p.a.cols * 16 <= 65536 && !env_int("GS_STREAM_NO_PRE", 0);
    if (pre) fn = pick_stream_pre(depth, shape);
    if (do_rainbow) {
        const uint32_t sc = pre ? 16u : 1u;
        if (p.a.col_is_u16) e = p.a.val_is_f16 ? rainbow_stream<uint16_t, __half>(p, pl, cta_row, sc, s) : rainbow_stream<uint16_t, float>(p, pl, cta_row, sc, s);
        else e = p.a.val_is_f16 ? rainbow_stream<uint32_t, __half>(p, pl, cta_row, sc, s) : rainbow_stream<uint32_t, float>(p, pl, cta_row, sc, s);
        if (e != cudaSuccess) return e;
    }
    if ((e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn)) != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, W * 32, dyn);
    if (e != cudaSuccess) return e;
    if (occ < ctas) return cudaErrorNotSupported;
    pl->fn = fn, pl->stream = true, pl->grid = n_cta, pl->threads = W * 32;
    pl->bulk_panel = env_int("GS_STREAM_BULK", 1) ? 1u : 0u;
    pl->pf = (uint32_t)env_int("GS_STREAM_PF", 2);
    snprintf(pl->name, sizeof pl->name, "k_b200_stream<N%d,%s,%s,d%d,%dx%dw%s,pdl>", p.N, p.a.col_is_u16 ? "u16" : "u32", p.a.val_is_f16 ? "f16" : "f32", depth, ctas, W,
             pre ? ",+rainbow,pre16" : do_rainbow ? ",+rainbow" : "");
    return cudaSuccess;
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

    // stream schedule: rows padded to 8, panel in shared memory, rows long enough for 256-wide steps to pay
    const char *sched = getenv("GS_ROW_SCHED");
    const bool long_rows = pl->n_rows > 0 && p.a.nnz / pl->n_rows >= 64;
    if (pl->vec == 8 && pl->smem_b && (sched ? strcmp(sched, "stream") == 0 : long_rows)) {
        cudaError_t es = stream_prepare(p, sm_count, pl, s);
        if (es == cudaSuccess) {
            *out = pl;
            return cudaSuccess;
        }
        if (es != cudaErrorNotSupported) return fail(es);
        cudaGetLastError();
        cudaFree(pl->col2), cudaFree(pl->val2), cudaFree(pl->w_nz), cudaFree(pl->w_row);
        pl->col2 = pl->val2 = nullptr, pl->w_nz = pl->w_row = nullptr;
    }
    // short rows and a panel that does not fit shared memory (C1: 37 non-zeros per row, 1.3 MB of B): a group of lanes per row, several
    // rows in flight per warp (b200_row.cuh, k_b200_row_short).  GS_ROW_GROUP = lanes per row (4 / 8 / 16), 0 = off.
    {
        const double mean_len = pl->n_rows ? (double)p.a.nnz / pl->n_rows : 0.0;
        (void)mean_len;
        // opt-in (GS_ROW_GROUP = 4 / 8 / 16): on the C1 stand-in every setting measured slower than one warp per row (80 / 49 / 33 us against 26.7): the
        // longest row (935 non-zeros) is walked by G lanes only, and the gathers of B are bound by L1 wavefronts either way (profiles/r2_c1_summary.md);
        // the range kernel's chunk variant (kernels_merge.cu) is the answer for short rows
        const int G = env_int("GS_ROW_GROUP", 0);
        if (!pl->smem_b && G > 0 && (!sched || strcmp(sched, "short") == 0)) {
            const void *sf = p.a.col_is_u16 ? (p.a.val_is_f16 ? pick_short<uint16_t, __half>(p.N, G) : pick_short<uint16_t, float>(p.N, G))
                                            : (p.a.val_is_f16 ? pick_short<uint32_t, __half>(p.N, G) : pick_short<uint32_t, float>(p.N, G));
            int occ = 0;
            if (sf && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sf, 256, 0) == cudaSuccess && occ >= 1) {
                pl->fn = sf, pl->threads = 256, pl->smem_bytes = 0;
                const uint64_t want = ((uint64_t)pl->n_rows * G + 255) / 256; // one row per group is enough CTAs; more than fit at once only queue
                pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)sm_count * occ, want));
                snprintf(pl->name, sizeof pl->name, "k_b200_row_short<N%d,%s,%s,%d lanes per row,%d rows in flight per warp,%dx%d CTAs>", p.N, p.a.col_is_u16 ? "u16" : "u32",
                         p.a.val_is_f16 ? "f16" : "f32", G, 32 / G, sm_count, occ);
                *out = pl;
                return cudaSuccess;
            }
            cudaGetLastError();
        }
    }
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
    if (pl->stream) {
        row::stream_args a;
        a.col = pl->col2 ? pl->col2 : p.a.col, a.val = pl->val2 ? pl->val2 : p.a.val, a.B = p.B, a.C = p.C;
        a.rowptr = pl->rowptr, a.rowid = pl->rowid, a.perm = p.a.perm;
        a.w_nz = pl->w_nz, a.w_row = pl->w_row;
        a.n_rows = pl->n_rows, a.begin_row = p.a.begin_row, a.b_bytes = (uint32_t)pl->smem_bytes, a.bulk_panel = pl->bulk_panel, a.pf = pl->pf;
        void *args[] = {&a};
        // programmatic dependent launch: only behind another launch of this kernel family on the same stream (the
        // kernel reads A and the panel before griddepcontrol.wait, so whatever ran before must not have written them)
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3(pl->grid), cfg.blockDim = dim3(pl->threads), cfg.dynamicSmemBytes = pl->smem_bytes + pl->table_bytes, cfg.stream = s;
        cfg.attrs = attr, cfg.numAttrs = (p.pdl && !p.dep_wait) ? 1 : 0; // the panel of B is staged before griddepcontrol.wait
        return cudaLaunchKernelExC(&cfg, pl->fn, args);
    }
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
    if (pl->w_nz) cudaFree(pl->w_nz);
    if (pl->w_row) cudaFree(pl->w_row);
    delete pl;
}
const char *b200_row_kernel_name(const b200_row_plan *pl) { return pl ? pl->name : ""; }

} // namespace gsd
