// Host side of the B200 tile kernel (GS_FAMILY_B200_TCGEN05): builds the tile stream of a row-direction chain once per
// plan (host/tile_format.hpp), owns the device copies, the stream-K workspace and the launch.  Device code and the
// design notes: b200_tile.cuh.
#include "b200_tile.cuh"
#include "../host/tile_format.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace gsd {

struct b200_tile_plan {
    uint32_t *tile_ptr = nullptr, *out_row = nullptr, *flags = nullptr; // flags: grid ready + grid slot_free + n_panels panel_seq
    uint16_t *off = nullptr, *val = nullptr;
    float *ws = nullptr;
    uint32_t n_tiles = 0, k_tiles = 0, n_panels = 0, grid = 0, ptr_cap = 0, seq = 0;
    uint64_t pairs = 0, nnz = 0;
    size_t smem = 0;
    const void *fn = nullptr;
    int N = 0;
    char name[96] = {0};
    CUtensorMap b_map;             // the dense matrix as TMA boxes of 64 k rows x one swizzle atom of columns
    const void *b_mapped = nullptr; // the device pointer b_map was encoded for
    uint32_t K = 0;
    unsigned long long *dbg = nullptr; // GS_TILE_DEBUG: DBG_RING launches x grid x 8 stamps x 2
};
constexpr int DBG_RING = 16;

namespace {

template <int N> void pick(bool dbg, const void *&fn, size_t &stage_bytes, int &stages, int &ctas)
{
    fn = dbg ? (const void *)tile::k_b200_tile<N, true> : (const void *)tile::k_b200_tile<N, false>;
    stages = tile::b_geom<N>::STAGES, ctas = tile::b_geom<N>::CTAS;
    stage_bytes = (size_t)tile::b_geom<N>::RING_BYTES;
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
cudaError_t encode_b_map(b200_tile_plan *pl, const void *B)
{
    encode_tiled_fn enc = encoder();
    if (!enc) return cudaErrorNotSupported;
    const int atom = pl->N < 64 ? pl->N : 64;
    cuuint64_t dims[2] = {(cuuint64_t)pl->N, (cuuint64_t)pl->K}, strides[1] = {(cuuint64_t)pl->N * 2};
    cuuint32_t box[2] = {(cuuint32_t)atom, (cuuint32_t)tile::TK}, estr[2] = {1, 1};
    CUtensorMapSwizzle sw = atom == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : atom == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    CUresult r = enc(&pl->b_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(B), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    pl->b_mapped = B;
    return cudaSuccess;
}

int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

} // namespace

cudaError_t b200_tile_prepare(const launch_params &p, int sm_count, b200_tile_plan **out, cudaStream_t s)
{
    *out = nullptr;
    if (!p.a.val_is_f16) return cudaErrorInvalidValue; // kind::f16 operands; fp32 values stay on the gather kernels (tf32 would break the 1e-5 bound)
    const uint32_t *rowptr = nullptr, *rowid = nullptr;
    uint32_t n_rows = 0;
    if (p.a.t_first_nz && p.a.t_first_row && !p.a.t_nz_relative) rowptr = p.a.t_first_nz, rowid = p.a.t_first_row, n_rows = p.a.n_t;
    else if (p.a.w_first_nz && p.a.w_first_row) rowptr = p.a.w_first_nz, rowid = p.a.w_first_row, n_rows = p.a.n_w;
    else if (p.a.tb_first_nz && p.a.tb_first_row) rowptr = p.a.tb_first_nz, rowid = p.a.tb_first_row, n_rows = p.a.n_tb;
    else return cudaErrorInvalidValue;
    b200_tile_plan *pl = new (std::nothrow) b200_tile_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    auto fail = [&](cudaError_t e) {
        b200_tile_free(pl);
        return e;
    };
    size_t stage_bytes = 0;
    int stages = 0, ctas = 1;
    const bool dbg = env_int("GS_TILE_DEBUG", 0) || env_int("GS_TILE_SKIP", 0);
    switch (p.N) {
    case 16: pick<16>(dbg, pl->fn, stage_bytes, stages, ctas); break;
    case 32: pick<32>(dbg, pl->fn, stage_bytes, stages, ctas); break;
    case 64: pick<64>(dbg, pl->fn, stage_bytes, stages, ctas); break;
    case 128: pick<128>(dbg, pl->fn, stage_bytes, stages, ctas); break;
    case 256: pick<256>(dbg, pl->fn, stage_bytes, stages, ctas); break;
    default: return fail(cudaErrorInvalidValue); // M=128 tcgen05.mma wants N % 16 == 0; N=8 stays on the stream kernel
    }
    pl->N = p.N, pl->K = p.a.cols;
    if (!encoder()) return fail(cudaErrorNotSupported);

    // the chain's arrays back to the host, once
    std::vector<uint32_t> rp(n_rows + 1), rid(n_rows), perm, colw(p.a.nnz);
    std::vector<uint16_t> val(p.a.nnz);
    cudaError_t e;
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

    gs::tile_stream ts;
    try {
        gs::build_tile_stream(rp.data(), colw.data(), val.data(), n_rows, p.a.cols, env_int("GS_TILE_NO_BANK_ORDER", 0) ? 0u : (uint32_t)tile::PRODUCER_WARPS, ts);
    } catch (...) {
        return fail(cudaErrorInvalidValue);
    }
    pl->n_panels = ts.n_panels, pl->k_tiles = ts.k_tiles, pl->n_tiles = ts.n_panels * ts.k_tiles;
    pl->pairs = ts.off.size(), pl->nnz = ts.nnz;
    pl->grid = (uint32_t)std::min<uint64_t>((uint64_t)env_int("GS_TILE_CTAS", sm_count), pl->n_tiles);
    if (pl->grid == 0 || (int)pl->grid > sm_count) return fail(cudaErrorInvalidValue); // the split-panel hand-off needs every CTA resident

    std::vector<uint32_t> orow((size_t)ts.n_panels * gs::TILE_ROWS, 0xffffffffu);
    for (uint32_t i = 0; i < n_rows; i++) {
        const uint32_t r = rid[i];
        if (p.a.perm && r >= perm.size()) return fail(cudaErrorInvalidValue);
        orow[i] = (p.a.perm ? perm[r] : r) + p.a.begin_row;
    }

    // the CTA's slice of tile_ptr sits behind the stages in shared memory
    pl->ptr_cap = (pl->n_tiles + pl->grid - 1) / pl->grid + 2;
    pl->smem = stage_bytes + 1024 + (size_t)pl->ptr_cap * 4;
    if (pl->smem > (size_t)(227 * 1024) / ctas - 2048) return fail(cudaErrorInvalidValue); // > ~2000 tiles per CTA: not a weight-sized matrix

    auto up = [&](void **dst, const void *src, size_t bytes) {
        cudaError_t e2 = cudaMalloc(dst, bytes + 16);
        if (e2 != cudaSuccess) return e2;
        return bytes ? cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, s) : cudaSuccess;
    };
    if ((e = up((void **)&pl->tile_ptr, ts.tile_ptr.data(), ts.tile_ptr.size() * 4)) != cudaSuccess) return fail(e);
    if ((e = up((void **)&pl->off, ts.off.data(), ts.off.size() * 2)) != cudaSuccess) return fail(e);
    if ((e = up((void **)&pl->val, ts.val.data(), ts.val.size() * 2)) != cudaSuccess) return fail(e);
    if ((e = up((void **)&pl->out_row, orow.data(), orow.size() * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc((void **)&pl->ws, (size_t)pl->grid * tile::TM * p.N * 4)) != cudaSuccess) return fail(e);
    const size_t n_flags = (size_t)pl->grid * 2 + pl->n_panels;
    if ((e = cudaMalloc((void **)&pl->flags, n_flags * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMemsetAsync(pl->flags, 0, n_flags * 4, s)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
    if ((e = cudaFuncSetAttribute(pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem)) != cudaSuccess) return fail(e);
    if ((e = cudaFuncSetAttribute(pl->fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)) != cudaSuccess) return fail(e);
    int occ = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pl->fn, tile::THREADS, pl->smem)) != cudaSuccess) return fail(e);
    if (occ < 1) return fail(cudaErrorInvalidConfiguration);
    snprintf(pl->name, sizeof pl->name, "k_b200_tile<N%d,f16,tcgen05 m128n%dk16,%d stages,streamK x%u,%d CTA/SM%s,pdl>", p.N, p.N, stages, pl->grid, occ,
             env_int("GS_TILE_NO_BANK_ORDER", 0) ? "" : ",bank-ordered");
    if (env_int("GS_TILE_DEBUG", 0)) {
        if ((e = cudaMalloc((void **)&pl->dbg, (size_t)DBG_RING * (pl->grid * 32 + 256) * 8)) != cudaSuccess) return fail(e);
        cudaMemset(pl->dbg, 0, (size_t)DBG_RING * (pl->grid * 32 + 256) * 8);
    }
    *out = pl;
    return cudaSuccess;
}

// GS_TILE_DEBUG: per-CTA time stamps of the last DBG_RING launches, relative to the first CTA start of the oldest one
static void dump_debug(b200_tile_plan *pl, cudaStream_t s)
{
    cudaStreamSynchronize(s);
    const size_t per = (size_t)pl->grid * 32 + 256;
    std::vector<unsigned long long> h((size_t)DBG_RING * per);
    cudaMemcpy(h.data(), pl->dbg, h.size() * 8, cudaMemcpyDeviceToHost);
    static const char *ev[16] = {"entry", "setup done", "first tile stored", "first full seen", "last mma issued", "producer fence (counter)", "all joined", "last acc ready",
                                 "producer wait zeroed", "producer store", "producer wait loads", "producer sts", "mma wait full", "mma wait acc_empty+bfull", "ahead wait empty", "ahead zero+arrive"};
    unsigned long long base = ~0ull;
    for (int L = 0; L < DBG_RING; L++)
        for (uint32_t c = 0; c < pl->grid; c++) {
            unsigned long long g = h[(size_t)L * per + ((size_t)c * 16 + 0) * 2];
            if (g && g < base) base = g;
        }
    for (int L = DBG_RING - 2; L < DBG_RING; L++) {
        const uint32_t slot = (pl->seq - (DBG_RING - 1 - L)) % DBG_RING;
        fprintf(stderr, "[tile debug] launch seq %u\n", pl->seq - (DBG_RING - 1 - L));
        for (int e = 0; e < 16; e++) {
            std::vector<double> t, ck;
            for (uint32_t c = 0; c < pl->grid; c++) {
                const size_t i = (size_t)slot * per + ((size_t)c * 16 + e) * 2, i0 = (size_t)slot * per + ((size_t)c * 16 + 0) * 2;
                if (!h[i]) continue;
                t.push_back(e < 8 && e != 5 ? (double)(h[i] - base) * 1e-3 : 0.0), ck.push_back(e < 8 && e != 5 ? (double)(h[i + 1] - h[i0 + 1]) : (double)h[i + 1]);
            }
            if (t.empty()) continue;
            std::sort(t.begin(), t.end()), std::sort(ck.begin(), ck.end());
            if (e == 0 && L == DBG_RING - 1) {
                const unsigned long long *tr = &h[(size_t)slot * per + (size_t)pl->grid * 32];
                const unsigned long long c0 = h[(size_t)slot * per + 1]; // CTA 0 entry clock
                fprintf(stderr, "  CTA 0 per-tile trace (cycles since entry): tile | zeroed | warp0 stored | warp9 stored | mma sees full | mma issued\n");
                for (int it = 0; it < 37; it++)
                    fprintf(stderr, "    %2d %8lld %8lld %8lld %8lld %8lld\n", it, (long long)(tr[it] - c0), (long long)(tr[40 + it] - c0), (long long)(tr[80 + it] - c0), (long long)(tr[120 + it] - c0),
                            (long long)(tr[160 + it] - c0));
            }
            fprintf(stderr, "  %-18s globaltimer us min %9.2f med %9.2f max %9.2f | cycles since entry min %8.0f med %8.0f max %8.0f (n=%zu)\n", ev[e], t.front(), t[t.size() / 2],
                    t.back(), ck.front(), ck[ck.size() / 2], ck.back(), t.size());
        }
    }
}

cudaError_t b200_tile_launch(b200_tile_plan *pl, const launch_params &p, cudaStream_t s)
{
    tile::tile_args a;
    a.tile_ptr = pl->tile_ptr, a.off = pl->off, a.val = pl->val, a.B = p.B, a.C = p.C, a.out_row = pl->out_row;
    a.ws = pl->ws, a.flags = pl->flags, a.slot_free = pl->flags + pl->grid, a.panel_seq = pl->flags + 2 * (size_t)pl->grid;
    a.n_tiles = pl->n_tiles, a.k_tiles = pl->k_tiles, a.K = p.a.cols, a.seq = pl->seq + 1, a.ptr_cap = pl->ptr_cap;
    a.skip = (uint32_t)env_int("GS_TILE_SKIP", 0);
    a.consumer_fence = (uint32_t)env_int("GS_TILE_CONSUMER_FENCE", 1);
    a.bulk_zero = (uint32_t)env_int("GS_TILE_BULK_ZERO", 0);
    a.dep_wait = p.dep_wait ? 1u : 0u;
    a.dbg = pl->dbg ? pl->dbg + (size_t)(a.seq % DBG_RING) * ((size_t)pl->grid * 32 + 256) : nullptr;
    if (pl->b_mapped != p.B) {
        cudaError_t em = encode_b_map(pl, p.B);
        if (em != cudaSuccess) return em;
    }
    void *args[] = {&a, &pl->b_map};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(pl->grid), cfg.blockDim = dim3(tile::THREADS), cfg.dynamicSmemBytes = pl->smem, cfg.stream = s;
    cfg.attrs = attr, cfg.numAttrs = p.pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelExC(&cfg, pl->fn, args);
    if (e == cudaSuccess) pl->seq = a.seq; // committed only now: a number no kernel publishes would make every later launch spin for ever
    if (pl->dbg && pl->seq == 100) dump_debug(pl, s);
    return e;
}

void b200_tile_free(b200_tile_plan *pl)
{
    if (!pl) return;
    cudaFree(pl->tile_ptr), cudaFree(pl->out_row), cudaFree(pl->flags), cudaFree(pl->off), cudaFree(pl->val), cudaFree(pl->ws), cudaFree(pl->dbg);
    delete pl;
}

const char *b200_tile_kernel_name(const b200_tile_plan *pl) { return pl ? pl->name : ""; }

// bytes one launch must move: the pair stream, the run starts, the output row table, B and C once each
uint64_t b200_tile_bytes(const b200_tile_plan *pl, uint64_t rows, uint64_t cols)
{
    return pl->pairs * 4 + ((uint64_t)pl->n_tiles + 1) * 4 + (uint64_t)pl->n_panels * tile::TM * 4 + cols * pl->N * 2 + rows * pl->N * 2;
}
uint64_t b200_tile_pairs(const b200_tile_plan *pl) { return pl->pairs; }

} // namespace gsd

// ---------------------------------------------------------------------------------------------------------------------
// gs_launch_no_alloc: the same kernel for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The stream-K workspace and the hand-off flags, which b200_tile_prepare allocates per
// plan, live here in one file-scope device array of a fixed, modest size that every launch through this entry point shares
// (every GPU that loads the library holds it until exit).  The caller passes B, C and the stream; nothing on this path
// allocates.  No counterpart in the reference.
namespace gsd {
namespace {
constexpr uint32_t NA_MAX_CTAS = 160;      // CTAs of one launch (one per SM; a B200 has 148)
constexpr uint32_t NA_MAX_N = 64;          // dense width
constexpr uint32_t NA_MAX_PANELS = 16384;  // row panels of 128 rows: 2 097 152 rows
constexpr size_t NA_WS_BYTES = (size_t)NA_MAX_CTAS * tile::TM * NA_MAX_N * sizeof(float); // 5 242 880
constexpr size_t NA_FLAG_WORDS = 2 * (size_t)NA_MAX_CTAS + NA_MAX_PANELS;                 // ready + slot_free + panel_seq
constexpr size_t NA_BYTES = NA_WS_BYTES + NA_FLAG_WORDS * sizeof(uint32_t);               // 5 309 696 bytes
constexpr int NA_MAX_DEVICES = 64;
} // namespace

// In host code this name is a host-side stand-in for the symbol, NOT a device address: it must never be handed to a kernel
// or to cudaMemsetAsync.  The device address comes from cudaGetSymbolAddress, once per device (no_alloc_scratch below).
static __device__ __align__(256) unsigned char g_no_alloc_scratch[NA_BYTES];

// device address of g_no_alloc_scratch on `device` (which must be current); looked up once per device
static cudaError_t no_alloc_scratch(int device, unsigned char **out)
{
    static unsigned char *addr[NA_MAX_DEVICES] = {nullptr};
    if (device < 0 || device >= NA_MAX_DEVICES) return cudaErrorInvalidDevice;
    unsigned char *p = __atomic_load_n(&addr[device], __ATOMIC_ACQUIRE);
    if (!p) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void *sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_no_alloc_scratch);
        if (e != cudaSuccess) return e;
        if (!sym || ((uintptr_t)sym & 255u)) return cudaErrorInvalidSymbol;
        p = (unsigned char *)sym;
        __atomic_store_n(&addr[device], p, __ATOMIC_RELEASE); // two threads racing here store the same address
    }
    *out = p;
    return cudaSuccess;
}

size_t b200_tile_no_alloc_scratch_bytes() { return NA_BYTES; }

// One launch of plan `pl` with the shared scratch.  `device` is the plan's device and is current.  *why names the limit
// that was exceeded when the return value is cudaErrorInvalidValue.  Calls that overlap on the GPU (two streams, or two
// host threads) race on the scratch; calls on one stream are ordered by the stream.
cudaError_t b200_tile_launch_no_alloc(const b200_tile_plan *pl, int device, const void *B, void *C, cudaStream_t s, const char **why)
{
    *why = nullptr;
    if (pl->N > (int)NA_MAX_N) {
        *why = "dense width above the no-alloc maximum of 64";
        return cudaErrorInvalidValue;
    }
    if (pl->grid > NA_MAX_CTAS) {
        *why = "more CTAs per launch than the no-alloc maximum of 160";
        return cudaErrorInvalidValue;
    }
    if (pl->n_panels > NA_MAX_PANELS) {
        *why = "more row panels than the no-alloc maximum of 16384 (2 097 152 rows)";
        return cudaErrorInvalidValue;
    }
    unsigned char *scratch = nullptr;
    cudaError_t e = no_alloc_scratch(device, &scratch);
    if (e != cudaSuccess) return e;
    float *ws = (float *)scratch;
    uint32_t *flags = (uint32_t *)(scratch + NA_WS_BYTES);
    // the array keeps the last call's contents: the flags compare against the launch number, so they start at zero
    // and this launch is number 1; the workspace is written before it is read and needs no clearing
    const size_t n_flags = 2 * (size_t)pl->grid + pl->n_panels;
    if ((e = cudaMemsetAsync(flags, 0, n_flags * sizeof(uint32_t), s)) != cudaSuccess) return e;
    tile::tile_args a;
    a.tile_ptr = pl->tile_ptr, a.off = pl->off, a.val = pl->val, a.B = B, a.C = C, a.out_row = pl->out_row;
    a.ws = ws, a.flags = flags, a.slot_free = flags + pl->grid, a.panel_seq = flags + 2 * (size_t)pl->grid;
    a.n_tiles = pl->n_tiles, a.k_tiles = pl->k_tiles, a.K = pl->K, a.seq = 1, a.ptr_cap = pl->ptr_cap;
    a.skip = 0, a.consumer_fence = 1, a.bulk_zero = 0, a.dep_wait = 0, a.dbg = nullptr;
    // the tensor map of B is a kernel parameter, encoded on the host into a local (the plan's own map stays as it is)
    b200_tile_plan tmp;
    tmp.N = pl->N, tmp.K = pl->K;
    if ((e = encode_b_map(&tmp, B)) != cudaSuccess) return e;
    void *args[] = {&a, &tmp.b_map};
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(pl->grid), cfg.blockDim = dim3(tile::THREADS), cfg.dynamicSmemBytes = pl->smem, cfg.stream = s;
    cfg.attrs = nullptr, cfg.numAttrs = 0; // no programmatic dependent launch: the memset above must have completed
    return cudaLaunchKernelExC(&cfg, pl->fn, args);
}

} // namespace gsd
