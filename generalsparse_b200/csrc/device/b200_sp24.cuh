// Device code of the B200 2:4 kernel (GS_FAMILY_B200_SP24): the sparse tensor-core path, tcgen05.mma.sp with the
// accumulator and the 2:4 metadata in tensor memory.  Host side: kernels_sp24.cu; the image: host/sp24_format.hpp.
//
// Nothing is decompressed here.  The host lays every 128 x 128 tile out as the bytes the tensor core reads -- 16 KB of
// compressed operand in the K-major 128-byte-swizzled layout, 2 KB of metadata in `tcgen05.cp.128x128b` order -- so a
// tile travels HBM -> (one bulk copy) -> shared memory -> tensor core without touching a register: 18 432 bytes per
// 16 384 matrix elements, 1.125 bytes per element, against 2 bytes dense and 4 bytes per non-zero in the
// (column id, value) streams of the gather kernels.  The kernel is a copy pipeline and is bound by HBM.
//   warp 0      one lane: bulk L2 prefetch of the CTA's (contiguous) run of tiles, and per tile, as soon as the tensor
//               core has released the stage, one cp.async.bulk of the tile image and the TMA boxes of the 128 x N
//               piece of B (row-major in global memory = the MN-major operand layout; the tensor map swizzles),
//               all completing on full[stage];
//   warp 1      one lane: tcgen05.cp of the stage's metadata into its 4 columns of tensor memory, 4 tcgen05.mma.sp
//               (M=128, N, K=32) per tile, commit to empty[stage], and to acc_full[] at the end of a row panel;
//   warps 4-7   epilogue: tcgen05.ld the accumulator (lane = row), convert, store the rows of C.
// A CTA may take P = 2 row panels at once (the unit of work is then P panels x one 128-column step, one bulk copy of P
// images): the 128 x N piece of B, which every tile of every panel needs again, is then fetched from L2 and written
// to shared memory once per 256 rows instead of once per 128 -- at N >= 64 B's re-reads, not A's stream, fill the path
// from L2 to the SMs (N=128: 32 KB of B per 18 KB of A).
// With remainder warps and gathers from L2 (RW = 8, TL = 0, N <= 16; the fallback of the tile-local form below): what the 2:4
// pattern cannot hold is multiplied inside the same CTA.  The CTA that
// owns a panel also owns its remainder rows; warps 8-15 (a pair of threads per row: two halves of the row's entries at
// N=8, the two 16-byte halves of B's row at N=16) gather B from L2 while the tile loop runs, leave the fp32 product in
// shared memory, and the epilogue adds it before the single rounding.  No second launch, no global hand-off.
// Tile-local remainder (TL, N <= 32): the remainder is cut along the tiles and travels with them; warps 8-15 (a pair of threads
// per tile row, accumulators in registers over the tiles of a segment) multiply a tile's entries with the 128 rows of B
// that the tile's own multiplication has brought to shared memory -- no gather leaves the SM, the work is split exactly
// like the tensor core's -- and hand the fp32 product of every segment to the epilogue through shared memory.
// Work split, split-panel hand-off and the ordering between overlapped launches (panel_seq / slot_free flags, no
// griddepcontrol.wait) are those of the tile kernel, b200_tile.cuh.
#pragma once
#include "b200_tile.cuh"

namespace gsd {
namespace sp24 {

using namespace tile; // barrier, TMA, tcgen05 and flag helpers

constexpr int SM_ROWS = 128, SK = 128, IMG_A = SM_ROWS * (SK / 2) * 2, IMG_E = SM_ROWS * 16, IMG_BYTES = IMG_A + IMG_E;
constexpr int PRODUCER_WARP = 0, ISSUE_WARP = 1, EPILOGUE_WARP0 = 4, SP_THREADS = 8 * 32, PF_AHEAD = 4;

struct sp24_args {
    const unsigned char *img; // n_units x P x IMG_BYTES (unit = P consecutive panels x one k step)
    void *C;
    const uint32_t *out_row; // n_panels * 128 output rows of C, 0xffffffff = no row
    const float *c_add;      // n_panels * 128 x N fp32 added to the accumulator before rounding (the remainder's product), or null
    const uint32_t *rem_seq; // per 32 rows: remainder CTAs that have delivered so far (rem_per_flag per launch, monotonic)
    uint32_t rem_per_flag;
    float *ws;               // grid x P x 128 x N fp32 partial panels
    uint32_t *flags;         // grid: launch number whose partial sits in the CTA's slot
    uint32_t *slot_free;     // grid: launch number whose partial has been consumed from the CTA's slot
    uint32_t *panel_seq;     // n_panels: launch number that last wrote the panel's rows of C
    uint32_t n_tiles, k_tiles, seq; // n_tiles: units
    uint32_t pf_ahead;              // units the bulk L2 prefetch runs ahead of the copy ring, 0 = no prefetch
    // fused remainder (kernels instantiated with remainder warps): CSR over the image's rows, one packed word per entry
    // (column id << 16 | fp16 value bits), multiplied by warps 8-15 of the CTA that owns the panel while the tile loop runs
    const uint32_t *rem_rowptr, *rem_pack;
    const void *B;
    uint32_t n_rows;
    // tile-local remainder (kernels instantiated with TL): every tile image is followed by the remainder entries of its 128 x 128 window
    // (host/sp24_format.hpp, sp24_tiled); one bulk copy of unit_bytes brings both, B's 128 rows sit b_off bytes into the stage
    uint32_t unit_bytes, b_off, stage_bytes, rem_on;
    uint32_t dep_wait; // 1: B may come from the kernel before this one on the stream: griddepcontrol.wait before B's first slice is requested
    // image without its empty tiles (group = 1): the units of panel p are panel_ptr[p] .. panel_ptr[p+1] (every panel keeps at least
    // one), unit u multiplies columns 128 unit_kt[u] ...; null = every tile present, unit t is tile (t / k_tiles, t % k_tiles)
    const uint32_t *panel_ptr, *unit_kt;
    uint32_t n_panels;
    uint32_t *dbg; // probe: CTA 0 leaves the 4 metadata columns of stage 0 (128 lanes x 4 words) here, or null
};

// where unit t lies: its panel and the panel's first and one-past-last unit
template <bool CP> __device__ __forceinline__ void seg_locate(const sp24_args &a, uint32_t t, uint32_t &panel, uint32_t &pbeg, uint32_t &pend)
{
    if constexpr (CP) {
        uint32_t lo = 0, hi = a.n_panels; // the largest p with panel_ptr[p] <= t (no panel is empty)
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (a.panel_ptr[mid] <= t) lo = mid;
            else hi = mid;
        }
        panel = lo, pbeg = a.panel_ptr[lo], pend = a.panel_ptr[lo + 1];
    } else panel = t / a.k_tiles, pbeg = panel * a.k_tiles, pend = pbeg + a.k_tiles;
}
// the segment that starts at unit t: kernels over the full image compute it from t (nothing kept between segments), kernels over
// the image without its empty tiles (CP) walk the panel table
template <bool CP> __device__ __forceinline__ void seg_begin(const sp24_args &a, uint32_t t, bool first, uint32_t &panel, uint32_t &pbeg, uint32_t &pend)
{
    if constexpr (CP) {
        if (first) seg_locate<true>(a, t, panel, pbeg, pend);
        else panel++, pbeg = pend, pend = a.panel_ptr[panel + 1];
    } else seg_locate<false>(a, t, panel, pbeg, pend);
}

__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// 128 rows x 16 bytes of shared memory -> 4 columns of tensor memory (lane = row)
__device__ __forceinline__ void tc_cp_128x128b(uint32_t taddr, uint64_t sdesc) { asm volatile("tcgen05.cp.cta_group::1.128x128b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory"); }
__device__ __forceinline__ void tc_mma_sp_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t e_tmem, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\ttcgen05.mma.sp.cta_group::1.kind::f16 [%0], %1, %2, [%3], %4, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc),
                 "r"(e_tmem), "r"(idesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void tc_ld4(uint32_t taddr, uint32_t (&r)[4])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// CH accumulator columns of the thread's row, issued only: tc_wait_ld() before the registers are read
template <int CH> __device__ __forceinline__ void tc_ld_issue(uint32_t taddr, uint32_t (&r)[CH])
{
    static_assert(CH == 8 || CH == 32, "chunk");
    if constexpr (CH == 8) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "r"(taddr));
    } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]),
                       "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
                       "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr));
    }
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// acc += a * b, a and b fp16 bits: the product is exact in fp32, one rounding (FHFMA)
__device__ __forceinline__ void fma_f16(float &acc, unsigned short a, unsigned short b) { asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(acc) : "h"(a), "h"(b)); }
template <int CH> __device__ __forceinline__ void add4(uint32_t (&r)[CH], const float4 (&x)[CH / 4])
{
#pragma unroll
    for (int j = 0; j < CH / 4; j++) {
        r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + x[j].x), r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + x[j].y);
        r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + x[j].z), r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + x[j].w);
    }
}

constexpr int pow2_at_least(int x) { return x <= 32 ? 32 : x <= 64 ? 64 : x <= 128 ? 128 : x <= 256 ? 256 : 512; }

// geometry for dense width N: the B operand is the 128 x N piece of the dense matrix, N contiguous = MN-major
template <int N, int P> struct geom {
    static constexpr int ATOM_N = N < 64 ? N : 64;           // dense columns per swizzle atom
    static constexpr int RB = ATOM_N * 2;                    // bytes of one k row inside an atom: 32 / 64 / 128
    static constexpr uint32_t LAYOUT = RB == 128 ? 2u : RB == 64 ? 4u : 6u; // SWIZZLE_128B / 64B / 32B
    static constexpr int GROUPS = N / ATOM_N;                // atoms along N
    static constexpr int GROUP_BYTES = SK * RB;              // one atom column: 128 k rows
    static constexpr int B_BYTES = GROUPS * GROUP_BYTES;     // = 128 * N * 2
    static constexpr int UNIT_BYTES = P * IMG_BYTES;         // the P tile images of a unit lie back to back
    static constexpr int STAGE_BYTES = UNIT_BYTES + B_BYTES; // multiples of 1024: the swizzle atoms stay aligned
    static constexpr int SBO = 8 * RB;                       // 8 k rows
    static constexpr int K32_STEP = 32 * RB;                 // bytes between the K=32 slices of one tile
    // two CTAs per SM (two launches in flight) while 2 x (stages + TMEM columns) fit
    static constexpr int CTAS = (P == 1 && N <= 128) ? 2 : 1;
    static constexpr int STAGES = P == 1 ? (N <= 32 ? 4 : N == 64 ? 3 : 2) : (N <= 64 ? 4 : 3);
    static constexpr int ACC = P * N <= 64 ? 2 : 1;         // accumulator sets in tensor memory (segments in flight), P accumulators each
    static constexpr int ECOL0 = ACC * P * N;                // the metadata columns sit behind the accumulators
    static constexpr int TMEM_COLS = pow2_at_least(ECOL0 + 4 * P * STAGES);
    static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
    // accumulator columns the epilogue moves per step: with one CTA per SM nothing hides the epilogue of a launch's last
    // panel, so its loads (tensor memory, the other CTAs' partials, the remainder's product) go out 32 columns at a time;
    // with two CTAs per SM the next launch covers it and the registers stay under the cap that lets two CTAs fit
    static constexpr int CH = (CTAS == 1 && N >= 32) ? 32 : 8;
    static constexpr int MIN_CTAS = CTAS == 1 ? 1 : 4;       // __launch_bounds__: 4 x 256 threads = 64 registers, room for the remainder pass beside two CTAs
    // instruction descriptor (UMMA::InstrDescriptor): sparse, D fp32, A/B fp16, A K-major, B MN-major, N >> 3, M >> 4
    static constexpr uint32_t IDESC = (1u << 2) | (1u << 4) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(SM_ROWS >> 4) << 24);
    static_assert(TMEM_COLS * CTAS <= 512 && ECOL0 + 4 * P * STAGES <= 512, "tensor memory");
    static_assert(RING_BYTES + 2048 <= 227 * 1024 / CTAS, "shared memory");
    static_assert(STAGE_BYTES % 1024 == 0, "stage alignment");
};

// N: the width the tensor core multiplies (a multiple of 16); NL <= N: the width of B and C in memory (NL = 8 runs as N = 16,
// the tensor map returns zeros for the columns B does not have)
// TL: 0 = no tile-local remainder, else the stages of the copy ring (4 where two CTAs with four stages fit an SM, else 3)
template <int N, int P, int NL, int RW, int TL = 0, bool CP = false> __global__ void __launch_bounds__(SP_THREADS + 32 * RW, (RW ? 2 : geom<N, P>::MIN_CTAS)) k_b200_sp24(sp24_args a, const __grid_constant__ CUtensorMap b_map)
{
    using G = geom<N, P>;
    constexpr int S = TL ? TL : G::STAGES, ACC = G::ACC, CH = G::CH;
    static_assert(NL % CH == 0, "chunk");
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    static_assert(RW == 0 || (RW == 8 && NL <= (TL ? 64 : 16) && P == 1 && G::CTAS == 2), "remainder warps");
    static_assert(!CP || P == 1, "the image without its empty tiles has one panel per unit");
    static_assert(TL == 0 || (TL >= 2 && TL <= 4 && RW == 8), "the tile-local remainder is multiplied by the remainder warps");
    // bytes of a unit in global memory (= one bulk copy), offset of B's rows inside a stage, stage stride
    const uint32_t UNIT = TL ? a.unit_bytes : (uint32_t)G::UNIT_BYTES, BOFF = TL ? a.b_off : (uint32_t)G::UNIT_BYTES, STG = TL ? a.stage_bytes : (uint32_t)G::STAGE_BYTES;
    __shared__ __align__(8) uint64_t bars[2 * S + 8];
    __shared__ uint32_t tmem_slot;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u; // 1024-byte alignment of the stages (128-byte swizzle atoms)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t n_cta = gridDim.x, cta = blockIdx.x;
    const uint32_t t0 = (uint32_t)((uint64_t)cta * a.n_tiles / n_cta), t1 = (uint32_t)((uint64_t)(cta + 1) * a.n_tiles / n_cta);
    const uint32_t n_local = t1 - t0;
    const uint32_t KT = a.k_tiles;

    const uint32_t bar0 = smem_u32(bars);
    // full[s] = bar0 + 8 s, empty[s] = + 8 (S + s), acc_full[b] = + 8 (2 S + b), acc_empty[b] = + 8 (2 S + 2 + b)
    const uint32_t full0 = bar0, empty0 = bar0 + 8u * S, accf0 = bar0 + 16u * S, acce0 = accf0 + 16u;
    // rem_full[b] = acce0 + 16 + 8 b, rem_empty[b] = + 16: the two remainder buffers behind the ring (128 rows x NL fp32 each)
    const uint32_t remf0 = acce0 + 16u, reme0 = remf0 + 16u;
    float *rem_buf = (float *)(smem_raw + (smem0 - smem_u32(smem_raw)) + S * STG);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) mbar_init(&bars[s], 1), mbar_init(&bars[S + s], TL ? 1 + RW : 1); // TL: the remainder warps read the stage too
        mbar_init(&bars[2 * S], 1), mbar_init(&bars[2 * S + 1], 1);
        mbar_init(&bars[2 * S + 2], 4), mbar_init(&bars[2 * S + 3], 4);
        if (RW) mbar_init(&bars[2 * S + 4], RW), mbar_init(&bars[2 * S + 5], RW), mbar_init(&bars[2 * S + 6], 4), mbar_init(&bars[2 * S + 7], 4);
    }
    if (warp == ISSUE_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"((uint32_t)G::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    row::pdl_launch_dependents();

    if (warp == PRODUCER_WARP) {
        // ------------------------------------------------------------------------------------------ copies
        if (lane == 0) {
            const unsigned char *img = a.img + (size_t)t0 * UNIT;
            const uint32_t pf = a.pf_ahead ? (uint32_t)S + a.pf_ahead : 0u;
            for (uint32_t it = 0; it < pf && it < n_local; it++) l2_prefetch_bulk(img + (size_t)it * UNIT, UNIT);
            uint32_t s = 0, ph = 0, kt = CP ? a.unit_kt[t0] : t0 % KT, it0 = 0;
            if (a.dep_wait) {
                // B may be what the previous kernel on the stream is still writing (a dependent chain: step i+1 multiplies step i's C).  The
                // launch was allowed to start early all the same: the images of A for the whole ring go out now, griddepcontrol.wait then
                // holds this lane until the predecessor has completed and flushed, and only then are B's slices requested.  Prologue,
                // tensor-memory allocation, L2 prefetch and the first S images overlap the predecessor's tail instead of following it.
                const uint32_t n0 = min((uint32_t)S, n_local);
                for (uint32_t it = 0; it < n0; it++) {
                    mbar_expect_tx(full0 + 8u * it, UNIT + (uint32_t)G::B_BYTES);
                    bulk_load(smem0 + it * STG, img + (size_t)it * UNIT, UNIT, full0 + 8u * it);
                }
                row::pdl_wait();
                for (uint32_t it = 0; it < n0; it++) {
                    uint32_t kt_next = 0;
                    if (CP && it + 1 < n_local) kt_next = a.unit_kt[t0 + it + 1];
#pragma unroll
                    for (int g = 0; g < G::GROUPS; g++) tma_load_2d(smem0 + it * STG + BOFF + g * G::GROUP_BYTES, &b_map, (uint32_t)(g * G::ATOM_N), kt * SK, full0 + 8u * it);
                    if (CP) kt = kt_next;
                    else if (++kt == KT) kt = 0;
                    if (++s == S) s = 0, ph ^= 1u;
                }
                it0 = n0;
            }
            for (uint32_t it = it0; it < n_local; it++) {
                uint32_t kt_next = 0; // requested before the wait below, so the table's latency hides behind it
                if (CP && it + 1 < n_local) kt_next = a.unit_kt[t0 + it + 1];
                if (pf && it + pf < n_local) l2_prefetch_bulk(img + (size_t)(it + pf) * UNIT, UNIT);
                mbar_wait(empty0 + 8u * s, ph ^ 1u); // the tensor core has finished reading the stage
                const uint32_t st = smem0 + s * STG, fb = full0 + 8u * s;
                mbar_expect_tx(fb, UNIT + (uint32_t)G::B_BYTES);
                bulk_load(st, img + (size_t)it * UNIT, UNIT, fb);
#pragma unroll
                for (int g = 0; g < G::GROUPS; g++) tma_load_2d(st + BOFF + g * G::GROUP_BYTES, &b_map, (uint32_t)(g * G::ATOM_N), kt * SK, fb); // rows past K come back as zeros
                if (CP) kt = kt_next;
                else if (++kt == KT) kt = 0;
                if (++s == S) s = 0, ph ^= 1u;
            }
        }
        __syncwarp();
    } else if (warp == ISSUE_WARP) {
        // ------------------------------------------------------------------------------------------ MMA issue
        if (lane == 0) {
            uint32_t it = 0, seg = 0, s = 0, ph = 0, panel = 0, pbeg = 0, pend = 0;
            const uint64_t adesc0 = smem_desc(smem0, 16, 1024, 2u), edesc0 = smem_desc(smem0 + IMG_A, 128, 128, 0u),
                           bdesc0 = smem_desc(smem0 + BOFF, G::GROUP_BYTES, G::SBO, G::LAYOUT);
            while (it < n_local) {
                seg_begin<CP>(a, t0 + it, it == 0, panel, pbeg, pend);
                const uint32_t len = min(pend - (t0 + it), n_local - it), ab = seg % ACC;
                mbar_wait(acce0 + 8u * ab, ((seg / ACC) & 1u) ^ 1u);
                tc_fence_after();
                for (uint32_t i = 0; i < len; i++) {
                    mbar_wait(full0 + 8u * s, ph);
                    tc_fence_after();
                    const uint64_t so = (uint64_t)(s * (STG >> 4));
#pragma unroll
                    for (uint32_t h = 0; h < (uint32_t)P; h++) {
                        const uint64_t ho = so + (uint64_t)(h * (IMG_BYTES >> 4));
                        const uint32_t ecol = (uint32_t)G::ECOL0 + 4u * (s * P + h), d = tmem_base + (ab * P + h) * N;
                        tc_cp_128x128b(tmem_base + ecol, edesc0 + ho);
#pragma unroll
                        for (uint32_t k = 0; k < SK / 32; k++) // the metadata column pair is addressed, the column inside it goes into the descriptor
                            tc_mma_sp_f16(d, adesc0 + ho + (uint64_t)(2 * k), bdesc0 + so + (uint64_t)(k * (G::K32_STEP >> 4)), tmem_base + ((ecol + k) & ~1u), G::IDESC | ((ecol + k) & 1u),
                                          (i > 0 || k > 0) ? 1u : 0u);
                    }
                    tc_commit(empty0 + 8u * s);
                    if (++s == S) s = 0, ph ^= 1u;
                }
                tc_commit(accf0 + 8u * ab);
                it += len, seg++;
            }
        }
        __syncwarp();
    } else if (warp >= EPILOGUE_WARP0 && warp < EPILOGUE_WARP0 + 4) {
        // ------------------------------------------------------------------------------------------ epilogue
        const uint32_t q = warp - EPILOGUE_WARP0, rip = q * 32 + lane, et = threadIdx.x - EPILOGUE_WARP0 * 32;
        __half *C = (__half *)a.C;
        uint32_t it = 0, seg = 0, own = 0; // own: owner segments so far (the remainder warps count the same way)
        uint32_t unit_panel = 0, pbeg = 0, pend = 0;
        auto wait_for = [&](const uint32_t *flag, uint32_t v) { // one lane polls, the warp follows
            if (lane == 0)
                while (ld_acquire(flag) != v) __nanosleep(32);
            __syncwarp();
        };
        while (it < n_local) {
            seg_begin<CP>(a, t0 + it, it == 0, unit_panel, pbeg, pend);
            const uint32_t t = t0 + it, len = min(pend - t, n_local - it), ab = seg % ACC;
            const bool partial_in = t > pbeg;                         // someone else owns the panels: leave partials
            const bool owner_wait = !partial_in && len < pend - pbeg; // this CTA owns the panels and others hold their tail
            // what does not depend on the accumulator goes first, while the tile loop is still running
            uint32_t orow[P], c_last = cta; // contributors: the CTAs after this one whose first unit lies inside the same panels
            if (partial_in) wait_for(a.slot_free + cta, a.seq - 1); // the previous launch's owner has read this CTA's slot
            else {
#pragma unroll
                for (int h = 0; h < P; h++) {
                    const uint32_t panel = unit_panel * P + h;
                    wait_for(a.panel_seq + panel, a.seq - 1); // the previous launch has written these rows of C
                    orow[h] = a.out_row[(size_t)panel * SM_ROWS + rip];
                    if (a.c_add) { // this launch's remainder product of the warp's 32 rows (a later launch's may already have replaced the flag)
                        if (lane == 0)
                            while ((int32_t)(ld_acquire(a.rem_seq + panel * 4u + q) - a.seq * a.rem_per_flag) < 0) __nanosleep(32);
                        __syncwarp();
                    }
                }
                if (owner_wait) {
                    const uint64_t panel_end = pend;
                    while (c_last + 1 < n_cta && (uint64_t)(c_last + 1) * a.n_tiles / n_cta < panel_end) c_last++;
                }
            }
            mbar_wait_idle(accf0 + 8u * ab, (seg / ACC) & 1u);
            tc_fence_after();
            if (a.dbg && cta == 0 && seg == 0) {
                uint32_t e[4];
                tc_ld4(tmem_base + ((q * 32u) << 16) + (uint32_t)G::ECOL0, e);
                for (int j = 0; j < 4; j++) a.dbg[rip * 4 + j] = e[j];
            }
            if (partial_in) {
#pragma unroll
                for (int h = 0; h < P; h++) {
                    const uint32_t taddr = tmem_base + ((q * 32u) << 16) + (ab * P + h) * N;
                    float *slot = a.ws + (((size_t)cta * P + h) * SM_ROWS + rip) * NL;
                    if constexpr (TL) mbar_wait(remf0, seg & 1u); // the segment's share of the remainder goes into the partial
#pragma unroll 1
                    for (int c0 = 0; c0 < NL; c0 += CH) {
                        uint32_t r[CH];
                        float4 xa[CH / 4];
                        tc_ld_issue<CH>(taddr + c0, r);
                        if constexpr (TL) {
#pragma unroll
                            for (int j = 0; j < CH / 4; j++) xa[j] = ((const float4 *)(rem_buf + rip * NL + c0))[j];
                        }
                        tc_wait_ld();
                        if constexpr (TL) add4<CH>(r, xa);
#pragma unroll
                        for (int j = 0; j < CH / 4; j++) __stcg((uint4 *)(slot + c0) + j, make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]));
                    }
                }
                if constexpr (TL) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(reme0);
                }
                __threadfence();
                bar_sync(3, 128);
                if (et == 0) st_release(a.flags + cta, a.seq);
            } else {
                for (uint32_t c2 = cta + 1; c2 <= c_last; c2++) wait_for(a.flags + c2, a.seq);
#pragma unroll
                for (int h = 0; h < P; h++) {
                    const uint32_t panel = unit_panel * P + h, taddr = tmem_base + ((q * 32u) << 16) + (ab * P + h) * N;
                    const float *add;
                    if constexpr (TL) { // the segment's remainder product, single buffer
                        add = rem_buf + rip * NL;
                        mbar_wait(remf0, seg & 1u);
                    } else if constexpr (RW != 0) { // the remainder warps of this CTA left the product of the panel's remainder rows in shared memory
                        add = a.rem_rowptr ? rem_buf + ((own & 1u) * SM_ROWS + rip) * NL : nullptr;
                        if (add) mbar_wait(remf0 + 8u * (own & 1u), (own >> 1) & 1u);
                    } else add = (a.c_add && orow[h] != 0xffffffffu) ? a.c_add + ((size_t)panel * SM_ROWS + rip) * NL : nullptr;
                    const bool h1 = c_last >= cta + 1, h2 = c_last >= cta + 2;
                    const float *s1 = a.ws + (((size_t)(cta + 1) * P + h) * SM_ROWS + rip) * NL, *s2 = a.ws + (((size_t)(cta + 2) * P + h) * SM_ROWS + rip) * NL;
#pragma unroll 1
                    for (int c0 = 0; c0 < NL; c0 += CH) { // every load of the step goes out before the first is used; the sums keep the order k, then the remainder
                        uint32_t r[CH];
                        float4 x1[CH / 4], x2[CH / 4], xa[CH / 4];
                        tc_ld_issue<CH>(taddr + c0, r);
                        if (h1) {
#pragma unroll
                            for (int j = 0; j < CH / 4; j++) x1[j] = __ldcg((const float4 *)(s1 + c0) + j);
                        }
                        if (h2) {
#pragma unroll
                            for (int j = 0; j < CH / 4; j++) x2[j] = __ldcg((const float4 *)(s2 + c0) + j);
                        }
                        if (add) {
#pragma unroll
                            for (int j = 0; j < CH / 4; j++) xa[j] = RW ? ((const float4 *)(add + c0))[j] : __ldcg((const float4 *)(add + c0) + j);
                        }
                        tc_wait_ld();
                        if (h1) add4<CH>(r, x1);
                        if (h2) add4<CH>(r, x2);
                        for (uint32_t c2 = cta + 3; c2 <= c_last; c2++) { // a panel spread over more than three CTAs: small matrices only
                            const float4 *slot = (const float4 *)(a.ws + (((size_t)c2 * P + h) * SM_ROWS + rip) * NL + c0);
#pragma unroll
                            for (int j = 0; j < CH / 4; j++) x1[j] = __ldcg(slot + j);
                            add4<CH>(r, x1);
                        }
                        if (add) add4<CH>(r, xa);
                        if (orow[h] != 0xffffffffu) {
#pragma unroll
                            for (int j8 = 0; j8 < CH / 8; j8++) {
                                uint32_t hh[4];
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    __half2 pr = __floats2half2_rn(__uint_as_float(r[8 * j8 + 2 * j]), __uint_as_float(r[8 * j8 + 2 * j + 1]));
                                    hh[j] = *(uint32_t *)&pr;
                                }
                                *(uint4 *)(C + (size_t)orow[h] * NL + c0 + 8 * j8) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
                            }
                        }
                    }
                }
                if constexpr (TL) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(reme0);
                } else if (RW != 0 && a.rem_rowptr) { // the remainder buffer is free again
                    __syncwarp();
                    if (lane == 0) mbar_arrive(reme0 + 8u * (own & 1u));
                    own++;
                }
                __threadfence();
                bar_sync(3, 128);
                if (et == 0) {
#pragma unroll
                    for (int h = 0; h < P; h++) st_release(a.panel_seq + unit_panel * P + h, a.seq);
                    for (uint32_t c2 = cta + 1; c2 <= c_last; c2++) st_release(a.slot_free + c2, a.seq);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acce0 + 8u * ab);
            it += len, seg++;
        }
    } else if (TL && warp >= EPILOGUE_WARP0 + 4) {
        // ------------------------------------------------------------------------------------------ remainder, tile-local
        // a pair of threads per tile row; NL = 8: the two take alternate entries, NL = 16: each takes a 16-byte half of B's
        // row, NL = 32: each takes 32 bytes.  B's rows lie as the tensor map swizzled them; an entry carries the byte offset of
        // its row's chunk 0 (host/sp24_format.hpp, sp24_b_row_offset), chunk c is at offset ^ (c << 4).  fp16 x fp16 + fp32 in
        // one instruction (fma.rn.f32.f16: the exact product, one rounding -- what fmaf of the converted operands gives).
        if constexpr (NL == 32) {
            // NL = 32 (round 2): FOUR threads per tile row, one 16-byte chunk of B's row each, the tile's rows in two passes of 64.  The pair-of-
            // threads form below (32 bytes per thread) issues 128-bit loads whose eight lanes per wavefront point at eight different rows of B:
            // ~960 shared-memory wavefronts per tile, 58 % of them bank conflicts -- more than the MMA operands and the copies together, and what
            // holds C2 N = 32 at 58 % of its roofline (profiles/traffic.json C2_N32_half_sp24; the remainder skipped: 9.95 us against 14.6).  With
            // four lanes on one row a wavefront's eight lanes read two whole 64-byte rows, which collide only when both fall into the same half
            // of the banks.  Same entries per row, same order of the sums; a warp now spans 8 rows instead of 16 (less waiting for its longest).
            const uint32_t rt = threadIdx.x - (EPILOGUE_WARP0 + 4) * 32, rq = rt >> 2, coff = (rt & 3u) << 4;
            const unsigned char *ring = smem_raw + (smem0 - smem_u32(smem_raw));
            uint32_t it = 0, seg = 0, s = 0, ph = 0, panel = 0, pbeg = 0, pend = 0;
            while (it < n_local) {
                seg_begin<CP>(a, t0 + it, it == 0, panel, pbeg, pend);
                const uint32_t len = min(pend - (t0 + it), n_local - it);
                float acc[2][8];
#pragma unroll
                for (int j = 0; j < 8; j++) acc[0][j] = acc[1][j] = 0.f;
                for (uint32_t i = 0; i < len; i++) {
                    mbar_wait(full0 + 8u * s, ph);
                    if (a.rem_on) {
                        const unsigned char *st = ring + s * STG, *bs = st + BOFF;
                        const unsigned short *start = (const unsigned short *)(st + IMG_BYTES);
                        const uint32_t *ent = (const uint32_t *)(st + IMG_BYTES + 272);
#pragma unroll
                        for (int pass = 0; pass < 2; pass++) {
                            const uint32_t rrow = rq + 64u * pass, e1 = start[rrow + 1];
                            uint32_t e = start[rrow];
                            // two entries per round; the next round's entry words are requested before this round's rows of B
                            uint32_t w0 = e < e1 ? ent[e] : 0u, w1 = e + 1 < e1 ? ent[e + 1] : 0u;
#pragma unroll 1
                            for (; e < e1; e += 2) {
                                const bool two = e + 1 < e1; // a missing second entry reads row 0 and is not added
                                const uint32_t n0 = e + 2 < e1 ? ent[e + 2] : 0u, n1 = e + 3 < e1 ? ent[e + 3] : 0u;
                                const uint4 b0 = *(const uint4 *)(bs + ((w0 >> 16) ^ coff)), b1 = *(const uint4 *)(bs + ((w1 >> 16) ^ coff));
                                const unsigned short v0 = (unsigned short)(w0 & 0xffffu), v1 = (unsigned short)(w1 & 0xffffu);
                                const uint32_t q0[4] = {b0.x, b0.y, b0.z, b0.w}, q1[4] = {b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                                for (int k = 0; k < 4; k++) fma_f16(acc[pass][2 * k], v0, (unsigned short)(q0[k] & 0xffffu)), fma_f16(acc[pass][2 * k + 1], v0, (unsigned short)(q0[k] >> 16));
                                if (two) {
#pragma unroll
                                    for (int k = 0; k < 4; k++) fma_f16(acc[pass][2 * k], v1, (unsigned short)(q1[k] & 0xffffu)), fma_f16(acc[pass][2 * k + 1], v1, (unsigned short)(q1[k] >> 16));
                                }
                                w0 = n0, w1 = n1;
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty0 + 8u * s); // this warp has read the stage
                    if (++s == S) s = 0, ph ^= 1u;
                }
                mbar_wait(reme0, (seg & 1u) ^ 1u); // the epilogue has read the previous segment's product
#pragma unroll
                for (int pass = 0; pass < 2; pass++) {
                    float4 *dst = (float4 *)(rem_buf + (rq + 64u * pass) * NL + (rt & 3u) * 8u);
                    dst[0] = make_float4(acc[pass][0], acc[pass][1], acc[pass][2], acc[pass][3]);
                    dst[1] = make_float4(acc[pass][4], acc[pass][5], acc[pass][6], acc[pass][7]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(remf0);
                it += len, seg++;
            }
        } else {
        constexpr int CPT = NL >= 32 ? NL / 16 : 1, NA = 8 * CPT; // NL = 64: 64 bytes of B's row per thread, one entry per round (registers)
        constexpr bool TWO = NL <= 32;
        const uint32_t rt = threadIdx.x - (EPILOGUE_WARP0 + 4) * 32, rrow = rt >> 1, sub = rt & 1u;
        const uint32_t estart = NL == 8 ? sub : 0u, estep = NL == 8 ? 2u : 1u, c16 = NL == 8 ? 0u : sub * CPT;
        const unsigned char *ring = smem_raw + (smem0 - smem_u32(smem_raw));
        uint32_t it = 0, seg = 0, s = 0, ph = 0, panel = 0, pbeg = 0, pend = 0;
        while (it < n_local) {
            seg_begin<CP>(a, t0 + it, it == 0, panel, pbeg, pend);
            const uint32_t len = min(pend - (t0 + it), n_local - it);
            float acc[NA];
#pragma unroll
            for (int j = 0; j < NA; j++) acc[j] = 0.f;
            for (uint32_t i = 0; i < len; i++) {
                mbar_wait(full0 + 8u * s, ph);
                if (a.rem_on) {
                    const unsigned char *st = ring + s * STG, *bs = st + BOFF;
                    const unsigned short *start = (const unsigned short *)(st + IMG_BYTES);
                    const uint32_t *ent = (const uint32_t *)(st + IMG_BYTES + 272);
                    const uint32_t e1 = start[rrow + 1];
                    uint32_t e = start[rrow] + estart;
                    // two entries per round; the next round's entry words are requested before this round's rows of B
                    constexpr uint32_t PER = TWO ? 2u : 1u; // entries per round
                    uint32_t w0 = e < e1 ? ent[e] : 0u, w1 = TWO && e + estep < e1 ? ent[e + estep] : 0u;
#pragma unroll 1
                    for (; e < e1; e += PER * estep) {
                        const bool two = TWO && e + estep < e1; // a missing second entry reads row 0 and is not added
                        const uint32_t n0 = e + PER * estep < e1 ? ent[e + PER * estep] : 0u, n1 = TWO && e + 3u * estep < e1 ? ent[e + 3u * estep] : 0u;
                        uint4 b0[CPT], b1[TWO ? CPT : 1];
#pragma unroll
                        for (int c = 0; c < CPT; c++) {
                            b0[c] = *(const uint4 *)(bs + ((w0 >> 16) ^ ((c16 + c) << 4)));
                            if constexpr (TWO) b1[c] = *(const uint4 *)(bs + ((w1 >> 16) ^ ((c16 + c) << 4)));
                        }
                        const unsigned short v0 = (unsigned short)(w0 & 0xffffu), v1 = (unsigned short)(w1 & 0xffffu);
#pragma unroll
                        for (int c = 0; c < CPT; c++) {
                            const uint32_t q0[4] = {b0[c].x, b0[c].y, b0[c].z, b0[c].w};
#pragma unroll
                            for (int k = 0; k < 4; k++) fma_f16(acc[8 * c + 2 * k], v0, (unsigned short)(q0[k] & 0xffffu)), fma_f16(acc[8 * c + 2 * k + 1], v0, (unsigned short)(q0[k] >> 16));
                            if constexpr (TWO) {
                                const uint32_t q1[4] = {b1[c].x, b1[c].y, b1[c].z, b1[c].w};
                                if (two) {
#pragma unroll
                                    for (int k = 0; k < 4; k++) fma_f16(acc[8 * c + 2 * k], v1, (unsigned short)(q1[k] & 0xffffu)), fma_f16(acc[8 * c + 2 * k + 1], v1, (unsigned short)(q1[k] >> 16));
                                }
                            }
                        }
                        w0 = n0, w1 = n1;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty0 + 8u * s); // this warp has read the stage
                if (++s == S) s = 0, ph ^= 1u;
            }
            mbar_wait(reme0, (seg & 1u) ^ 1u); // the epilogue has read the previous segment's product
            if (NL == 8) {
#pragma unroll
                for (int k = 0; k < 8; k++) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], 1);
            }
            if (NL != 8 || sub == 0u) {
                float4 *dst = (float4 *)(rem_buf + rrow * NL + (NL == 8 ? 0u : sub * NA));
#pragma unroll
                for (int j = 0; j < NA / 4; j++) dst[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(remf0);
            it += len, seg++;
        }
        } // NL != 32
    } else if (RW != 0 && warp >= EPILOGUE_WARP0 + 4) {
        // ------------------------------------------------------------------------------------------ remainder
        // a pair of threads per row of the panel this CTA owns: at NL = 8 the two take alternate entries of the row, at
        // NL = 16 both take every entry and each gathers its 16-byte half of B's row
        constexpr int LPR = NL / 8;
        const uint32_t rt = threadIdx.x - (EPILOGUE_WARP0 + 4) * 32, rrow = rt >> 1, sub = rt & 1u;
        const uint32_t slice = LPR == 2 ? sub : 0u, estart = LPR == 2 ? 0u : sub, estep = LPR == 2 ? 1u : 2u;
        const __half *B = (const __half *)a.B;
        uint32_t it = 0, own = 0, panel = 0, pbeg = 0, pend = 0;
        while (it < n_local && a.rem_rowptr) {
            seg_begin<CP>(a, t0 + it, it == 0, panel, pbeg, pend);
            const uint32_t t = t0 + it, len = min(pend - t, n_local - it);
            if (t == pbeg) {
                const uint32_t b = own & 1u;
                mbar_wait(reme0 + 8u * b, ((own >> 1) & 1u) ^ 1u); // the epilogue has read what this buffer held two panels back
                const uint32_t r = panel * SM_ROWS + rrow;
                float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (r < a.n_rows) {
                    const uint32_t e1 = a.rem_rowptr[r + 1];
#pragma unroll 1
                    for (uint32_t e = a.rem_rowptr[r] + estart; e < e1; e += 4u * estep) {
                        uint32_t w[4];
                        uint4 bb[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) w[j] = e + j * estep < e1 ? __ldg(a.rem_pack + e + j * estep) : 0xffffffffu;
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if (e + j * estep < e1) bb[j] = __ldg((const uint4 *)(B + (size_t)(w[j] >> 16) * NL + slice * 8u));
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if (e + j * estep < e1) {
                                const float v = __half2float(__ushort_as_half((unsigned short)(w[j] & 0xffffu)));
                                const uint32_t q[4] = {bb[j].x, bb[j].y, bb[j].z, bb[j].w};
#pragma unroll
                                for (int k = 0; k < 4; k++) {
                                    const float2 f = __half22float2(*(const __half2 *)&q[k]);
                                    acc[2 * k] = fmaf(v, f.x, acc[2 * k]), acc[2 * k + 1] = fmaf(v, f.y, acc[2 * k + 1]);
                                }
                            }
                    }
                }
                if (LPR == 1) {
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], 1);
                }
                if (LPR == 2 || sub == 0u) {
                    float4 *dst = (float4 *)(rem_buf + (b * SM_ROWS + rrow) * NL + slice * 8u);
                    dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]), dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(remf0 + 8u * b);
                own++;
            }
            it += len;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == ISSUE_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G::TMEM_COLS) : "memory");
}

} // namespace sp24
} // namespace gsd
