// Host side of the B200 2:4 kernel (GS_FAMILY_B200_SP24): builds the 2:4 image of a row-direction chain once per plan
// (host/sp24_format.hpp), owns the device copies, the stream-K workspace and the launch.  Device code: b200_sp24.cuh.
#include "b200_sp24.cuh"
#include "sp24_build.hpp"
#include "../host/sp24_format.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace gsd {

struct b200_sp24_plan {
    bool poisoned = false; // a launch failed half way (remainder pass out, 2:4 kernel not): the flags are out of step
    unsigned char *img = nullptr;
    uint32_t *out_row = nullptr, *flags = nullptr; // flags: grid ready + grid slot_free + n_panels panel_seq + rem_ctas rem_seq
    float *ws = nullptr, *c_add = nullptr;         // c_add: two buffers (launch parity) of n_panels * 128 x N fp32, when there is a remainder
    uint32_t *rem_rowptr = nullptr, *rem_col = nullptr, *rem_pack = nullptr; // rem_pack: column id << 16 | value bits, the fused pass's entry stream
    bool fused = false;                                                       // the remainder is multiplied by warps of the 2:4 kernel itself
    bool tiled = false;                                                       // ... from tile-local blocks that travel with the tile images (sp24_tiled)
    uint32_t unit_bytes = 0, b_off = 0, stage_bytes = 0;
    uint32_t *panel_ptr = nullptr, *unit_kt = nullptr; // the image without its empty tiles (b200_sp24.cuh), null = every tile present
    uint32_t n_units_full = 0;
    int threads = 0;
    uint16_t *rem_val = nullptr;
    uint32_t n_tiles = 0, k_tiles = 0, n_panels = 0, grid = 0, seq = 0, n_rows = 0, group = 1; // n_tiles: units of `group` panels x one k step
    uint64_t nnz = 0, rem_nnz = 0;
    size_t smem = 0;
    const void *fn = nullptr, *rem_fn = nullptr;
    int N = 0;
    char name[192] = {0};
    CUtensorMap b_map;              // the dense matrix as TMA boxes of 128 k rows x one swizzle atom of columns
    const void *b_mapped = nullptr; // the device pointer b_map was encoded for
    uint32_t K = 0;
    uint32_t *dbg = nullptr; // GS_SP24_PROBE: the metadata columns of CTA 0's first tile as tensor memory holds them
};

namespace sp24 {

// a remainder CTA is four warps and four rows: small enough (5 K registers, no shared memory) that six of them fit an SM beside two
// CTAs of the 2:4 kernel, so the pass runs under the previous launch's tile loop instead of waiting for its registers
constexpr int REM_THREADS = 128, REM_ROWS = 4, REM_PER_FLAG = 32 / REM_ROWS;

struct rem_args {
    const uint32_t *rowptr, *col;
    const uint16_t *val;
    const void *B;
    float *c_add;              // this launch's buffer
    uint32_t *rem_seq;         // per 32 rows (the rows of one epilogue warp of the 2:4 kernel): remainder CTAs that have delivered, REM_PER_FLAG per launch
    const uint32_t *panel_seq; // per panel: launch number that last wrote the panel's rows of C (the buffer's reader two launches back is done)
    uint32_t n_rows, seq;
};

// Product of the remainder (what the 2:4 pattern could not hold: 7.7 % of the non-zeros of a 70 %-sparse unstructured
// matrix, nothing for a 2:4-pruned one) with B, fp32, one row per warp.  32 (column id, value) entries of a row are one
// coalesced load per lane and are handed round by shuffles; a group of LPR = N/8 lanes gathers one row of B with
// 16-byte loads, the 32/LPR groups take the entries in turn, so every gather of a 32-entry block is independent of the
// others (two dependent memory round trips per block, not per entry); a butterfly adds the groups.
template <int N> __global__ void __launch_bounds__(REM_THREADS) k_b200_sp24_rem(rem_args a)
{
    constexpr int LPR = N / 8, GROUPS = 32 / LPR, RPW = REM_ROWS / (REM_THREADS / 32);
    static_assert(LPR >= 1 && LPR <= 32, "dense width");
    row::pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, grp = lane / LPR, sub = lane % LPR;
    const uint32_t panel = blockIdx.x / (128 / REM_ROWS);
    static_assert(RPW == 1, "one row per warp");
    if (threadIdx.x == 0 && a.seq >= 2)
        while ((int32_t)(tile::ld_acquire(a.panel_seq + panel) - (a.seq - 2)) < 0) __nanosleep(32);
    __syncthreads();
    const __half *B = (const __half *)a.B;
#pragma unroll 1
    for (int i = 0; i < RPW; i++) {
        const uint32_t r = blockIdx.x * REM_ROWS + warp * RPW + i;
        if (r >= a.n_rows) break;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const uint32_t e0 = a.rowptr[r], e1 = a.rowptr[r + 1];
#pragma unroll 2
        for (uint32_t e = e0; e < e1; e += 32) {
            const bool have = e + lane < e1;
            const uint32_t mycol = have ? __ldg(a.col + e + lane) : 0u;
            const float myval = have ? __half2float(__ushort_as_half(__ldg(a.val + e + lane))) : 0.0f;
#pragma unroll
            for (int j = 0; j < LPR; j++) {
                const int src = j * GROUPS + grp;
                const uint32_t c = __shfl_sync(0xffffffffu, mycol, src);
                const float v = __shfl_sync(0xffffffffu, myval, src);
                if (e + src < e1) {
                    const uint4 b = __ldg((const uint4 *)(B + (size_t)c * N + sub * 8));
                    const uint32_t w[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float2 f = __half22float2(*(const __half2 *)&w[q]);
                        acc[2 * q] = fmaf(v, f.x, acc[2 * q]), acc[2 * q + 1] = fmaf(v, f.y, acc[2 * q + 1]);
                    }
                }
            }
        }
#pragma unroll
        for (int off = LPR; off < 32; off <<= 1)
#pragma unroll
            for (int j = 0; j < 8; j++) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], off);
        if (grp == 0) {
            float4 *o = (float4 *)(a.c_add + (size_t)r * N + sub * 8);
            __stcg(o, make_float4(acc[0], acc[1], acc[2], acc[3])), __stcg(o + 1, make_float4(acc[4], acc[5], acc[6], acc[7]));
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.rem_seq + blockIdx.x / REM_PER_FLAG) : "memory");
}

} // namespace sp24

namespace {

template <int N, int P, int NL = N, int RW = 0> void pick(const void *&fn, const void *&rem_fn, size_t &ring_bytes, int &stages, int &ctas)
{
    fn = (const void *)sp24::k_b200_sp24<N, P, NL, RW>, rem_fn = (const void *)sp24::k_b200_sp24_rem<NL>;
    stages = sp24::geom<N, P>::STAGES, ctas = sp24::geom<N, P>::CTAS, ring_bytes = (size_t)sp24::geom<N, P>::RING_BYTES;
}
// tile-local remainder: the stage holds the unit (image + remainder block, rounded up to 1 KB) and B's 128 rows
int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

constexpr size_t TILED_SMEM_MAX = (228 * 1024 - 2 * (1024 + 2048 + 64)) / 2; // two CTAs per SM: alignment slack, static and reserved shared memory
template <int N, int NL> void pick_tiled(b200_sp24_plan *pl, size_t &ring_bytes, int &stages)
{
    pl->b_off = (pl->unit_bytes + 1023u) & ~1023u, pl->stage_bytes = pl->b_off + (uint32_t)sp24::geom<N, 1>::B_BYTES;
    const size_t rem = (size_t)sp24::SM_ROWS * NL * 4; // the segment's remainder product
    stages = N == 64 ? 2 : (N <= 16 && 4 * (size_t)pl->stage_bytes + rem <= TILED_SMEM_MAX && env_int("GS_SP24_TL_STAGES", 4) >= 4) ? 4 : 3;
    if constexpr (N <= 16) pl->fn = stages == 4 ? (const void *)sp24::k_b200_sp24<N, 1, NL, 8, 4> : (const void *)sp24::k_b200_sp24<N, 1, NL, 8, 3>;
    else if constexpr (N == 64) pl->fn = (const void *)sp24::k_b200_sp24<N, 1, NL, 8, 2>; // two stages of 38 KB + 32 KB of product: two CTAs per SM
    else pl->fn = (const void *)sp24::k_b200_sp24<N, 1, NL, 8, 3>;
    ring_bytes = (size_t)stages * pl->stage_bytes + rem;
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime (libcuda is not linked: the library must load on a CPU-only box)
encode_tiled_fn encoder()
{
    static encode_tiled_fn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_fn)p;
    }();
    return fn;
}

// B is K x N fp16 row-major: dimension 0 = dense column (contiguous), dimension 1 = k
cudaError_t encode_b_map(b200_sp24_plan *pl, const void *B)
{
    encode_tiled_fn enc = encoder();
    if (!enc) return cudaErrorNotSupported;
    const int atom = pl->N < 16 ? 16 : pl->N < 64 ? pl->N : 64; // N=8 runs as N=16: the box is wider than the tensor, the rest comes back as zeros
    cuuint64_t dims[2] = {(cuuint64_t)pl->N, (cuuint64_t)pl->K}, strides[1] = {(cuuint64_t)pl->N * 2};
    cuuint32_t box[2] = {(cuuint32_t)atom, (cuuint32_t)sp24::SK}, estr[2] = {1, 1};
    CUtensorMapSwizzle sw = atom == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : atom == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    CUresult r = enc(&pl->b_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(B), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    pl->b_mapped = B;
    return cudaSuccess;
}


} // namespace

cudaError_t b200_sp24_prepare(const launch_params &p, int sm_count, bool allow_remainder, b200_sp24_plan **out, cudaStream_t s)
{
    *out = nullptr;
    if (!p.a.val_is_f16) return cudaErrorInvalidValue; // kind::f16 operands
    const uint32_t *rowptr = nullptr, *rowid = nullptr;
    uint32_t n_rows = 0;
    if (p.a.t_first_nz && p.a.t_first_row && !p.a.t_nz_relative) rowptr = p.a.t_first_nz, rowid = p.a.t_first_row, n_rows = p.a.n_t;
    else if (p.a.w_first_nz && p.a.w_first_row) rowptr = p.a.w_first_nz, rowid = p.a.w_first_row, n_rows = p.a.n_w;
    else if (p.a.tb_first_nz && p.a.tb_first_row) rowptr = p.a.tb_first_nz, rowid = p.a.tb_first_row, n_rows = p.a.n_tb;
    else return cudaErrorInvalidValue;
    b200_sp24_plan *pl = new (std::nothrow) b200_sp24_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    auto fail = [&](cudaError_t e) {
        b200_sp24_free(pl);
        return e;
    };
    size_t ring_bytes = 0;
    int stages = 0, ctas = 1;
    // row panels per unit of work: two halve the re-reads of B (b200_sp24.cuh) but measured slower at every N (profiles/r1_sp24_summary.md),
    // so one is the default; at N=256 two accumulators would fill tensor memory
    const int group = p.N == 256 ? 1 : std::max(1, std::min(2, env_int("GS_SP24_PANELS", 1)));
    pl->group = (uint32_t)group;
    switch (p.N * 4 + group) {
    case 8 * 4 + 1: pick<16, 1, 8>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 8 * 4 + 2: pick<16, 2, 8>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 16 * 4 + 1: pick<16, 1>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 32 * 4 + 1: pick<32, 1>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 64 * 4 + 1: pick<64, 1>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 128 * 4 + 1: pick<128, 1>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 256 * 4 + 1: pick<256, 1>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 16 * 4 + 2: pick<16, 2>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 32 * 4 + 2: pick<32, 2>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 64 * 4 + 2: pick<64, 2>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    case 128 * 4 + 2: pick<128, 2>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas); break;
    default: return fail(cudaErrorInvalidValue); // M=128 tcgen05.mma wants N % 16 == 0 (8 runs as 16)
    }
    pl->N = p.N, pl->K = p.a.cols, pl->n_rows = n_rows;
    if (!encoder()) return fail(cudaErrorNotSupported);

    // ---- the image built on the device (sp24_build.cu) wherever the default form applies: one panel per unit and a remainder that is either
    // empty or travels tile-local with the images.  The counts come first; the choices below are the host path's own rules on them.
    cudaError_t e;
    const int fuse = env_int("GS_SP24_FUSED_REM", 2);
    sp24_dev_image db;
    bool on_device = false;
    uint32_t dev_unit_bytes = 0;
    static thread_local bool force_host = false; // a device build that failed half-way (out of memory ...) is repeated on the host, once
    if (group == 1 && !force_host && env_int("GS_SP24_DEVICE_BUILD", 1)) {
        if (sp24_dev_count(p, rowptr, n_rows, s, db) == cudaSuccess && !db.bad) {
            if (db.rem_nnz == 0) on_device = true;
            else if (allow_remainder && fuse >= 2 && p.N <= 64 && db.max_entries <= 0xffffu) {
                dev_unit_bytes = (uint32_t)sp24::IMG_BYTES + gs::SP24_REM_HEADER + ((db.max_entries + 3u) & ~3u) * 4u;
                b200_sp24_plan probe = *pl; // what pick_tiled would choose, without committing to it
                probe.unit_bytes = dev_unit_bytes;
                size_t rb = 0;
                int st = 0;
                if (p.N == 8) pick_tiled<16, 8>(&probe, rb, st);
                else if (p.N == 16) pick_tiled<16, 16>(&probe, rb, st);
                else if (p.N == 32) pick_tiled<32, 32>(&probe, rb, st);
                else pick_tiled<64, 64>(&probe, rb, st);
                const double content = (double)db.T * sp24::IMG_BYTES + 4.0 * (double)db.rem_nnz + (double)gs::SP24_REM_HEADER * (double)db.T;
                on_device = rb <= TILED_SMEM_MAX && (double)db.T * dev_unit_bytes <= 1.10 * content; // else: owner-CTA gathers or a gather pass, built on the host
            }
        } else cudaGetLastError();
        if (!on_device) db.release();
    }
    auto fail_dev = [&](cudaError_t e2) {
        db.release();
        return fail(e2);
    };

    // the chain's arrays back to the host, once (host path)
    std::vector<uint32_t> rp, rid, perm, colw;
    std::vector<uint16_t> val;
    gs::sp24_image im;
    if (!on_device) {
    rp.resize(n_rows + 1), rid.resize(n_rows), colw.resize(p.a.nnz), val.resize(p.a.nnz);
    if ((e = cudaMemcpyAsync(rp.data(), rowptr, rp.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(rid.data(), rowid, rid.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(val.data(), p.a.val, val.size() * 2, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    if (p.a.perm) {
        perm.resize(p.a.rows);
        if ((e = cudaMemcpyAsync(perm.data(), p.a.perm, perm.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    }
    if (p.a.col_is_u16) {
        std::vector<uint16_t> c16(p.a.nnz);
        if ((e = cudaMemcpyAsync(c16.data(), p.a.col, c16.size() * 2, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        for (size_t i = 0; i < c16.size(); i++) colw[i] = c16[i];
    } else {
        if ((e = cudaMemcpyAsync(colw.data(), p.a.col, colw.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
    }
    if (rp[n_rows] > p.a.nnz) return fail(cudaErrorInvalidValue);

    try {
        gs::build_sp24_image(rp.data(), colw.data(), val.data(), n_rows, p.a.cols, (uint32_t)group, im);
    } catch (...) {
        return fail(cudaErrorInvalidValue);
    }
    } // host path
    const uint32_t im_n_panels = on_device ? db.n_panels : im.n_panels, im_k_tiles = on_device ? db.k_tiles : im.k_tiles;
    const uint64_t im_bytes = on_device ? (uint64_t)db.T * sp24::IMG_BYTES : im.bytes.size(), im_nnz = on_device ? db.img_nnz : im.nnz;
    pl->rem_nnz = on_device ? db.rem_nnz : im.rem_col.size();
    if (pl->rem_nnz && !allow_remainder) return fail_dev(cudaErrorInvalidValue); // not a 2:4 matrix
    // The remainder is multiplied inside the 2:4 kernel (b200_sp24.cuh), no second launch.  GS_SP24_FUSED_REM: 2 (default) = from
    // tile-local blocks at N <= 64, as long as two CTAs still fit an SM and the padding to the largest block costs < 10 % of the
    // stream; 1 = gathers by the CTA that owns the panel (N <= 16); 0 = a separate gather launch.
    gs::sp24_tiled tl;
    const uint32_t tl_row_bytes = p.N == 64 ? 128u : p.N == 32 ? 64u : 32u;
    if (pl->rem_nnz && fuse >= 2 && p.N <= 64 && group == 1) {
        if (on_device) tl.unit_bytes = dev_unit_bytes;
        else try {
            gs::localize_sp24_remainder(im, tl_row_bytes, tl);
        } catch (...) {
            tl = gs::sp24_tiled();
        }
        if (tl.unit_bytes) {
            const void *fn0 = pl->fn;
            const size_t ring0 = ring_bytes;
            const int stages0 = stages;
            pl->unit_bytes = tl.unit_bytes;
            if (p.N == 8) pick_tiled<16, 8>(pl, ring_bytes, stages);
            else if (p.N == 16) pick_tiled<16, 16>(pl, ring_bytes, stages);
            else if (p.N == 32) pick_tiled<32, 32>(pl, ring_bytes, stages);
            else pick_tiled<64, 64>(pl, ring_bytes, stages);
            const double content = (double)im_bytes + 4.0 * (double)pl->rem_nnz + (double)gs::SP24_REM_HEADER * (double)(im_bytes / gs::SP24_TILE_BYTES);
            const double tl_bytes = on_device ? (double)db.T * dev_unit_bytes : (double)tl.bytes.size();
            pl->tiled = ring_bytes <= TILED_SMEM_MAX && tl_bytes <= 1.10 * content;
            if (pl->tiled) ctas = 2;
            else pl->fn = fn0, ring_bytes = ring0, stages = stages0, pl->unit_bytes = 0;
        }
    }
    pl->fused = pl->tiled || (pl->rem_nnz && fuse >= 1 && p.N <= 16 && group == 1 && p.a.cols <= 65536);
    pl->threads = sp24::SP_THREADS + (pl->fused ? 8 * 32 : 0);
    if (pl->fused && !pl->tiled) {
        if (p.N == 8) pick<16, 1, 8, 8>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas);
        else pick<16, 1, 16, 8>(pl->fn, pl->rem_fn, ring_bytes, stages, ctas);
        ring_bytes += 2 * sp24::SM_ROWS * p.N * 4; // the two remainder buffers behind the ring
    }
    pl->n_panels = im_n_panels, pl->k_tiles = im_k_tiles, pl->n_tiles = im_n_panels / group * im_k_tiles, pl->nnz = im_nnz;
    pl->n_units_full = pl->n_tiles;
    // Tiles that hold nothing (banded, block-sparse matrices) are dropped from the stream when that saves at least a tenth of it; every
    // panel keeps one tile so that its rows of C are written.  GS_SP24_SKIP_EMPTY=0 keeps them all.
    gs::sp24_compact cp;
    const void *dev_compact = nullptr; // device path: the kernel over a panel table, if the fill decides to drop the empty tiles
    if (group == 1 && env_int("GS_SP24_SKIP_EMPTY", 1)) {
        const void *cfn = nullptr; // the kernel over a panel table: same shape as the one picked above
        if (pl->tiled) {
            if (p.N == 8) cfn = stages == 4 ? (const void *)sp24::k_b200_sp24<16, 1, 8, 8, 4, true> : (const void *)sp24::k_b200_sp24<16, 1, 8, 8, 3, true>;
            else if (p.N == 16) cfn = stages == 4 ? (const void *)sp24::k_b200_sp24<16, 1, 16, 8, 4, true> : (const void *)sp24::k_b200_sp24<16, 1, 16, 8, 3, true>;
            else if (p.N == 32) cfn = (const void *)sp24::k_b200_sp24<32, 1, 32, 8, 3, true>;
            else cfn = (const void *)sp24::k_b200_sp24<64, 1, 64, 8, 2, true>;
        } else if (pl->fused) cfn = p.N == 8 ? (const void *)sp24::k_b200_sp24<16, 1, 8, 8, 0, true> : (const void *)sp24::k_b200_sp24<16, 1, 16, 8, 0, true>;
        else switch (p.N) {
            case 8: cfn = (const void *)sp24::k_b200_sp24<16, 1, 8, 0, 0, true>; break;
            case 16: cfn = (const void *)sp24::k_b200_sp24<16, 1, 16, 0, 0, true>; break;
            case 32: cfn = (const void *)sp24::k_b200_sp24<32, 1, 32, 0, 0, true>; break;
            case 64: cfn = (const void *)sp24::k_b200_sp24<64, 1, 64, 0, 0, true>; break;
            case 128: cfn = (const void *)sp24::k_b200_sp24<128, 1, 128, 0, 0, true>; break;
            case 256: cfn = (const void *)sp24::k_b200_sp24<256, 1, 256, 0, 0, true>; break;
            }
        if (on_device) dev_compact = cfn;
        else if (cfn && gs::compact_sp24_units(pl->tiled ? tl.bytes.data() : im.bytes.data(), pl->tiled ? tl.unit_bytes : (size_t)sp24::IMG_BYTES, pl->tiled, im.n_panels, im.k_tiles, 0.9, cp)) {
            pl->fn = cfn;
            pl->n_tiles = (uint32_t)cp.unit_kt.size();
        }
    }
    if (on_device) { // the fill: image, metadata, tile-local blocks, and the unit / panel tables when the empty tiles leave the stream
        uint32_t n_units = 0;
        if (env_int("GS_SP24_DEVICE_BUILD", 1) == 2) e = cudaErrorMemoryAllocation; // (test hook: the fill "fails", the fallback below runs)
        else e = sp24_dev_fill(p, rowptr, n_rows, pl->tiled ? tl.unit_bytes : (uint32_t)sp24::IMG_BYTES, tl_row_bytes, dev_compact != nullptr, 0.9, s, db, &pl->img, &pl->panel_ptr, &pl->unit_kt, &n_units);
        if (e == cudaSuccess) {
            if (pl->panel_ptr) pl->fn = dev_compact, pl->n_tiles = n_units;
            e = sp24_dev_out_row(p, rowid, n_rows, im_n_panels, s, db, &pl->out_row);
        }
        if (e != cudaSuccess) { // the host builder takes over
            cudaGetLastError();
            fail_dev(e);
            force_host = true;
            e = b200_sp24_prepare(p, sm_count, allow_remainder, out, s);
            force_host = false;
            return e;
        }
        db.release();
    }
    pl->grid = (uint32_t)std::min<uint64_t>((uint64_t)env_int("GS_TILE_CTAS", sm_count), pl->n_tiles);
    if (pl->grid == 0 || (int)pl->grid > sm_count * ctas) return fail(cudaErrorInvalidValue); // the split-panel hand-off needs every CTA resident

    std::vector<uint32_t> orow(on_device ? 0 : (size_t)im.n_panels * sp24::SM_ROWS, 0xffffffffu);
    for (uint32_t i = 0; i < n_rows && !on_device; i++) {
        const uint32_t r = rid[i];
        if (p.a.perm && r >= perm.size()) return fail(cudaErrorInvalidValue);
        orow[i] = (p.a.perm ? perm[r] : r) + p.a.begin_row;
    }
    pl->smem = ring_bytes + 1024;

    auto up = [&](void **dst, const void *src, size_t bytes) {
        cudaError_t e2 = cudaMalloc(dst, bytes + 16);
        if (e2 != cudaSuccess) return e2;
        return bytes ? cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, s) : cudaSuccess;
    };
    if (on_device) {
        // the image, its tables and the row table are in place
    } else if (!cp.bytes.empty()) {
        if ((e = up((void **)&pl->img, cp.bytes.data(), cp.bytes.size())) != cudaSuccess) return fail(e);
        if ((e = up((void **)&pl->panel_ptr, cp.panel_ptr.data(), cp.panel_ptr.size() * 4)) != cudaSuccess) return fail(e);
        if ((e = up((void **)&pl->unit_kt, cp.unit_kt.data(), cp.unit_kt.size() * 4)) != cudaSuccess) return fail(e);
    } else if ((e = pl->tiled ? up((void **)&pl->img, tl.bytes.data(), tl.bytes.size()) : up((void **)&pl->img, im.bytes.data(), im.bytes.size())) != cudaSuccess) return fail(e);
    if (!on_device && (e = up((void **)&pl->out_row, orow.data(), orow.size() * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc((void **)&pl->ws, (size_t)pl->grid * group * sp24::SM_ROWS * p.N * 4)) != cudaSuccess) return fail(e);
    const size_t n_flags = (size_t)pl->grid * 2 + pl->n_panels + (size_t)pl->n_panels * 4;
    if ((e = cudaMalloc((void **)&pl->flags, n_flags * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMemsetAsync(pl->flags, 0, n_flags * 4, s)) != cudaSuccess) return fail(e);
    if (pl->tiled) {
        // everything the kernel needs is in the units
    } else if (pl->fused) {
        std::vector<uint32_t> pack(im.rem_col.size());
        for (size_t i = 0; i < pack.size(); i++) pack[i] = (im.rem_col[i] << 16) | im.rem_val[i];
        if ((e = up((void **)&pl->rem_rowptr, im.rem_rowptr.data(), im.rem_rowptr.size() * 4)) != cudaSuccess) return fail(e);
        if ((e = up((void **)&pl->rem_pack, pack.data(), pack.size() * 4)) != cudaSuccess) return fail(e);
    } else if (pl->rem_nnz) {
        if ((e = up((void **)&pl->rem_rowptr, im.rem_rowptr.data(), im.rem_rowptr.size() * 4)) != cudaSuccess) return fail(e);
        if ((e = up((void **)&pl->rem_col, im.rem_col.data(), im.rem_col.size() * 4)) != cudaSuccess) return fail(e);
        if ((e = up((void **)&pl->rem_val, im.rem_val.data(), im.rem_val.size() * 2)) != cudaSuccess) return fail(e);
        const size_t cb = (size_t)pl->n_panels * sp24::SM_ROWS * p.N * 4;
        if ((e = cudaMalloc((void **)&pl->c_add, 2 * cb)) != cudaSuccess) return fail(e);
        if ((e = cudaMemsetAsync(pl->c_add, 0, 2 * cb, s)) != cudaSuccess) return fail(e);
    }
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
    if ((e = cudaFuncSetAttribute(pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem)) != cudaSuccess) return fail(e);
    if ((e = cudaFuncSetAttribute(pl->fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)) != cudaSuccess) return fail(e);
    // the same shared-memory / L1 split for the remainder pass: CTAs of kernels that want different splits cannot share an SM, and the
    // pass is meant to run beside the resident CTAs of the 2:4 kernel
    if ((e = cudaFuncSetAttribute(pl->rem_fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)) != cudaSuccess) return fail(e);
    int occ = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pl->fn, pl->threads, pl->smem)) != cudaSuccess) return fail(e);
    if (occ < 1) return fail(cudaErrorInvalidConfiguration);
    // 512-thread CTAs at 64 registers fill the register file exactly with two CTAs; the occupancy calculator answers 1 for them, the
    // hardware runs 2 (ncu: launch__occupancy_limit_registers = launch__occupancy_limit_shared_mem = 2, profiles/r1_sp24_summary.md)
    if (pl->fused && ctas == 2 && occ == 1 && 2 * (pl->smem + 2048) <= 228 * 1024) occ = 2;
    char rem[48] = "", skip[40] = "";
    if (pl->panel_ptr) snprintf(skip, sizeof skip, ",%u of %u tiles", pl->n_tiles, pl->n_units_full);
    if (pl->rem_nnz) snprintf(rem, sizeof rem, ",+remainder %.1f%% (%s)", 100.0 * (double)pl->rem_nnz / (double)(pl->rem_nnz + pl->nnz), pl->tiled ? "tile-local, 8 warps" : pl->fused ? "8 warps of the owner CTA" : "gather pass");
    snprintf(pl->name, sizeof pl->name, "k_b200_sp24<N%d,f16,tcgen05.mma.sp m128n%dk32,%d panel%s/CTA,%d stages,streamK x%u,%d CTA/SM,pdl%s%s>", p.N, p.N < 16 ? 16 : p.N, group, group > 1 ? "s" : "", stages, pl->grid, occ,
             rem, skip);
    if (env_int("GS_SP24_PROBE", 0)) {
        if ((e = cudaMalloc((void **)&pl->dbg, 128 * 4 * 4)) != cudaSuccess) return fail(e);
        cudaMemset(pl->dbg, 0xff, 128 * 4 * 4);
    }
    *out = pl;
    return cudaSuccess;
}

cudaError_t b200_sp24_launch(b200_sp24_plan *pl, const launch_params &p, cudaStream_t s)
{
    if (pl->b_mapped != p.B) {
        cudaError_t em = encode_b_map(pl, p.B);
        if (em != cudaSuccess) return em;
    }
    // the launch number is committed only after every launch of this call succeeded: the kernels of later launches wait for flags == seq - 1, and a
    // number no kernel ever publishes would make them spin for ever.  A call that launched its remainder pass but not the 2:4 kernel leaves
    // rem_seq ahead of panel_seq: the plan refuses further launches instead of deadlocking.
    if (pl->poisoned) return cudaErrorLaunchFailure;
    const uint32_t seq = pl->seq + 1;
    uint32_t *slot_free = pl->flags + pl->grid, *panel_seq = pl->flags + 2 * (size_t)pl->grid, *rem_seq = panel_seq + pl->n_panels;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg;
    float *c_add = nullptr;
    bool rem_launched = false;
    static const bool skip_rem = env_int("GS_SP24_SKIP_REM", 0) != 0; // timing experiments only: wrong results on a matrix with a remainder
    if (pl->rem_nnz && !skip_rem && !pl->fused) {
        c_add = pl->c_add + (size_t)(seq & 1u) * pl->n_panels * sp24::SM_ROWS * pl->N;
        sp24::rem_args r;
        r.rowptr = pl->rem_rowptr, r.col = pl->rem_col, r.val = pl->rem_val, r.B = p.B, r.c_add = c_add, r.rem_seq = rem_seq, r.panel_seq = panel_seq;
        r.n_rows = pl->n_rows, r.seq = seq;
        void *rargs[] = {&r};
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3(pl->n_panels * (128 / sp24::REM_ROWS)), cfg.blockDim = dim3(sp24::REM_THREADS), cfg.stream = s;
        cfg.attrs = attr, cfg.numAttrs = (p.pdl && !p.dep_wait) ? 1 : 0; // the gather pass reads B at once
        cudaError_t e = cudaLaunchKernelExC(&cfg, pl->rem_fn, rargs);
        if (e != cudaSuccess) return e; // nothing launched: seq stays
        rem_launched = true;
    }
    sp24::sp24_args a;
    a.img = pl->img, a.C = p.C, a.out_row = pl->out_row, a.c_add = c_add, a.rem_seq = rem_seq, a.rem_per_flag = sp24::REM_PER_FLAG;
    a.ws = pl->ws, a.flags = pl->flags, a.slot_free = slot_free, a.panel_seq = panel_seq;
    a.n_tiles = pl->n_tiles, a.k_tiles = pl->k_tiles, a.seq = seq;
    static const uint32_t pf_ahead = (uint32_t)env_int("GS_SP24_PREFETCH", sp24::PF_AHEAD);
    a.pf_ahead = pf_ahead;
    a.dbg = pl->dbg;
    a.rem_rowptr = (pl->fused && !pl->tiled && !skip_rem) ? pl->rem_rowptr : nullptr, a.rem_pack = pl->rem_pack, a.B = p.B, a.n_rows = pl->n_rows;
    a.unit_bytes = pl->unit_bytes, a.b_off = pl->b_off, a.stage_bytes = pl->stage_bytes, a.rem_on = skip_rem ? 0u : 1u;
    a.panel_ptr = pl->panel_ptr, a.unit_kt = pl->unit_kt, a.n_panels = pl->n_panels;
    // B read only through the copy lane's TMA (no remainder, or the tile-local one, whose entries read B's rows from the stage): the launch
    // may start under programmatic dependent launch even when B belongs to the previous kernel -- the lane waits before B's first slice
    const bool can_wait = pl->tiled || !pl->rem_nnz || skip_rem;
    a.dep_wait = (p.dep_wait && can_wait) ? 1u : 0u;
    const bool pdl = p.pdl && (!p.dep_wait || can_wait);
    void *args[] = {&a, &pl->b_map};
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(pl->grid), cfg.blockDim = dim3(pl->threads), cfg.dynamicSmemBytes = pl->smem, cfg.stream = s;
    cfg.attrs = attr, cfg.numAttrs = (pdl || (pl->rem_nnz && !pl->fused)) ? 1 : 0; // behind its own remainder launch the 2:4 kernel always overlaps: the flags order them
    const cudaError_t e = cudaLaunchKernelExC(&cfg, pl->fn, args);
    if (e == cudaSuccess) pl->seq = seq;
    else if (rem_launched) pl->poisoned = true;
    return e;
}

void b200_sp24_free(b200_sp24_plan *pl)
{
    if (!pl) return;
    cudaFree(pl->img), cudaFree(pl->out_row), cudaFree(pl->flags), cudaFree(pl->ws), cudaFree(pl->c_add), cudaFree(pl->rem_rowptr), cudaFree(pl->rem_col), cudaFree(pl->rem_val), cudaFree(pl->rem_pack),
        cudaFree(pl->dbg), cudaFree(pl->panel_ptr), cudaFree(pl->unit_kt);
    delete pl;
}

const char *b200_sp24_kernel_name(const b200_sp24_plan *pl) { return pl ? pl->name : ""; }

// bytes one launch must move: the image, the remainder's streams, the output row table, B and C once each
uint64_t b200_sp24_bytes(const b200_sp24_plan *pl, uint64_t rows, uint64_t cols)
{
    if (pl->tiled) // the images, 4 bytes per remainder entry, the row starts of every block -- not the padding to the largest block
        return (uint64_t)pl->n_tiles * (sp24::IMG_BYTES + gs::SP24_REM_HEADER) + pl->rem_nnz * 4 + (uint64_t)pl->n_panels * sp24::SM_ROWS * 4 + cols * pl->N * 2 + rows * pl->N * 2;
    return (uint64_t)pl->n_tiles * pl->group * sp24::IMG_BYTES + pl->rem_nnz * (pl->fused ? 4 : 6) + (pl->rem_nnz ? ((uint64_t)pl->n_rows + 1) * 4 : 0) + (uint64_t)pl->n_panels * sp24::SM_ROWS * 4 + cols * pl->N * 2 +
           rows * pl->N * 2;
}
uint64_t b200_sp24_remainder(const b200_sp24_plan *pl) { return pl->rem_nnz; }
int b200_sp24_launches(const b200_sp24_plan *pl) { return pl->rem_nnz && !pl->fused ? 2 : 1; }
cudaError_t b200_sp24_probe(const b200_sp24_plan *pl, uint32_t *out512)
{
    if (!pl->dbg) return cudaErrorInvalidValue;
    return cudaMemcpy(out512, pl->dbg, 128 * 4 * 4, cudaMemcpyDeviceToHost);
}

} // namespace gsd
