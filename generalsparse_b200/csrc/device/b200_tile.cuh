// Device code of the B200 tile kernel (GS_FAMILY_B200_TCGEN05): sparse-tile decompress to shared memory + tcgen05.mma
// with the accumulator in tensor memory.  Host side and the format: kernels_tile.cu, host/tile_format.hpp.
//
// Why a tensor-core path at all: at N >= 16 the gather kernels (b200_stream.cuh) are bound by the shared-memory data
// pipe and the FMA pipe, not by HBM -- 13.5 M non-zeros x N=32 is 433 M fp32 FMAs and 433 MB of LDS traffic per launch,
// 12 us and 23 us on 148 SMs against 8.4 us of HBM time.  A 30 %-dense 128 x 64 tile rebuilt in shared memory costs one
// 2-byte store per non-zero whatever N is, and the contraction itself (2.9 GFLOP dense at N=32) is 2 us of tensor pipe.
//
// One CTA per SM and launch, 16 warps x 64 registers, sized so that TWO CTAs fit an SM: with programmatic dependent launch the next
// launch's CTAs become resident beside this launch's and their prologue, pipeline fill and first tiles overlap this
// launch's tail (the split-panel hand-off below), which a 37-tile CTA otherwise pays in full every launch.
//   warp 11     one lane runs ahead of everybody: bulk L2 prefetch of the pair runs 8 tiles ahead (UBLKPF), and, as
//               soon as the tensor core has released a stage, one st.bulk that zeroes the 16 KB image of A's tile
//               (UMEMSETS) + an arrive on zeroed[stage], and the TMA load of the 64 x N piece of B (row-major in
//               global memory = the MN-major operand layout; the tensor map applies the 32/64/128-byte swizzle),
//               completing on full[stage];
//   warps 0-9   producers, every warp on every tile and independent of the others (no CTA barrier in the loop):
//               scatter the warp's share of the tile's (offset, value) pairs with 2-byte stores into the K-major 128-byte-swizzled image (128-bit
//               loads of 8 offsets + 8 values per lane; two register sets, so the loads of the next tile are in
//               flight while this tile is stored), fence to the async proxy, arrive on full[stage];
//   warp 10     one elected lane issues 4 tcgen05.mma (M=128, N, K=16) per tile, commits to empty[stage], and to
//               acc_full[] at the end of a row panel; two accumulators in TMEM alternate between panels;
//   warps 12-15 epilogue: tcgen05.ld the accumulator (lane = row), convert, store the rows of C.
// The tiles are dealt in equal contiguous runs to the CTAs (a run crosses row panels: stream-K).  A panel split over
// several CTAs is summed by the CTA that holds its first tile, in k order, from fp32 partials the others leave in a
// per-CTA workspace slot (release/acquire flag carrying the launch sequence number): no atomics, bit-reproducible.
// Launches of one plan overlap on the SMs, so nothing here waits for the previous grid as a whole
// (griddepcontrol.wait would gate every launch on the slowest CTA of the one before); the two hazards between
// launches are ordered point to point with the same kind of flag: a panel's rows of C are written by launch L only
// after launch L-1 has written them (panel_seq), and a workspace slot is refilled by launch L only after launch L-1's
// owner has consumed it (slot_free).
#pragma once
#include "b200_stream.cuh"

#include <cuda.h> // CUtensorMap (the type only: the encoder is reached through cudaGetDriverEntryPoint)

namespace gsd {
namespace tile {

using row::mbar_init;
using row::smem_u32;

constexpr int TM = 128, TK = 64, A_BYTES = TM * TK * 2;
constexpr int PRODUCER_WARPS = 10, PRODUCER_THREADS = PRODUCER_WARPS * 32, MMA_WARP = 10, AHEAD_WARP = 11, EPI_WARP0 = 12, THREADS = 16 * 32, PF_TILES = 8;
constexpr int PASS_PAIRS = PRODUCER_THREADS * 8; // pairs one sweep of the producer warps stores: 2560, above a 30 %-dense tile (2458 +- 41)
constexpr int DEPTH = 2;                         // register sets of pairs per producer lane: DEPTH - 1 tiles are in flight while one is stored

struct tile_args {
    const uint32_t *tile_ptr; // n_tiles + 1
    const uint16_t *off, *val;
    const void *B;
    void *C;
    const uint32_t *out_row; // n_panels * 128 output rows of C, 0xffffffff = no row
    float *ws;               // grid x 128 x N fp32 partial panels
    uint32_t *flags;         // grid: launch number whose partial sits in the CTA's slot
    uint32_t *slot_free;     // grid: launch number whose partial has been consumed from the CTA's slot
    uint32_t *panel_seq;     // n_panels: launch number that last wrote the panel's rows of C
    uint32_t n_tiles, k_tiles, K, seq, ptr_cap;
    uint32_t bulk_zero;      // 1: zero A's image with one st.bulk instead of the warp's STS.128
    uint32_t consumer_fence; // 1: the generic->async proxy fence runs on the mma lane after its acquire of full[], not in every producer lane
    uint32_t skip; // GS_TILE_SKIP (timing experiments, wrong results): 1 no scatter, 2 no zero fill, 4 no mma, 8 no B copy, 16 no pair loads, 32 no fence
    uint32_t dep_wait;       // 1: griddepcontrol.wait before B's first slice is requested (B may belong to the previous kernel on the stream)
    unsigned long long *dbg; // GS_TILE_DEBUG: 8 (globaltimer, clock64) stamps per CTA, or null
};

__device__ __forceinline__ void stamp(const tile_args &a, int ev)
{
    if (a.dbg) {
        unsigned long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        a.dbg[((size_t)blockIdx.x * 16 + ev) * 2] = g, a.dbg[((size_t)blockIdx.x * 16 + ev) * 2 + 1] = (unsigned long long)clock64();
    }
}

// per-tile trace of CTA 0: clock64 of role event `ev` (0 stage zeroed, 1 warp 0 stored, 2 warp 9 stored, 3 mma sees full, 4 mma issued) at tile `it`
__device__ __forceinline__ void trace(const tile_args &a, int ev, uint32_t it)
{
    if (a.dbg && blockIdx.x == 0 && it < 40) a.dbg[(size_t)gridDim.x * 32 + ev * 40 + it] = (unsigned long long)clock64();
}
__device__ __forceinline__ void counter(const tile_args &a, int ev, long long v)
{
    if (a.dbg) a.dbg[((size_t)blockIdx.x * 16 + ev) * 2] = 1, a.dbg[((size_t)blockIdx.x * 16 + ev) * 2 + 1] = (unsigned long long)v;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity);
// timed wait: adds the cycles spent to acc (debug builds of the timeline only cost a clock read per wait)
__device__ __forceinline__ void mbar_wait_t(uint32_t bar, uint32_t parity, long long &acc)
{
    long long t = clock64();
    mbar_wait(bar, parity);
    acc += clock64() - t;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar),
                 "r"(parity)
                 : "memory");
}
// a wait that is expected to last long (the epilogue warps sit out the whole tile loop): every retry of try_wait is an
// operation in the same memory-instruction queue the producers' stores go through, so poll rarely; the hardware
// still wakes the thread when the phase completes inside the suspend window
__device__ __forceinline__ void mbar_wait_idle(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar),
                 "r"(parity), "r"(100000u)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc),
                 "r"(idesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]),
                   "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t (&r)[8])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void sts16(uint32_t addr, unsigned short v) { asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(v) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) { asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ void smem_zero_bulk(uint32_t addr, uint32_t bytes) { asm volatile("st.bulk.weak.shared::cta [%0], %1, 0;" ::"r"(addr), "l"((uint64_t)bytes) : "memory"); }
__device__ __forceinline__ void l2_prefetch_bulk(const void *p, uint32_t bytes) { asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
// one box of the dense matrix (c0 = first dense column, c1 = first k row) into shared memory, swizzled by the tensor map
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t c0, uint32_t c1, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(uint32_t *p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// Shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp, UMMA::SmemDescriptor): start address, leading and
// stride byte offsets in 16-byte units, version 1, layout type in bits 61-63.
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout)
{
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46) |
           ((uint64_t)layout << 61);
}

// geometry of the B operand (64 x N piece of the dense matrix, N contiguous = MN-major) for dense width N
template <int N> struct b_geom {
    static constexpr int ATOM_N = N < 64 ? N : 64;           // dense columns per swizzle atom
    static constexpr int RB = ATOM_N * 2;                    // bytes of one k row inside an atom: 32 / 64 / 128
    static constexpr int SWZ_MASK = RB / 16 - 1;             // 16-byte chunks per row - 1
    static constexpr uint32_t LAYOUT = RB == 128 ? 2u : RB == 64 ? 4u : 6u; // SWIZZLE_128B / 64B / 32B
    static constexpr int GROUPS = N / ATOM_N;                // atoms along N
    static constexpr int GROUP_BYTES = TK * RB;              // one atom column: 64 k rows
    static constexpr int BYTES = GROUPS * GROUP_BYTES;       // = 64 * N * 2
    static constexpr int B_SLOT = BYTES < 1024 ? 1024 : BYTES;
    static constexpr int CHUNKS = BYTES / 16;
    static constexpr int SBO = 8 * RB;                       // 8 k rows
    static constexpr int K16_STEP = 16 * RB;                 // bytes between the K=16 slices of one tile
    static constexpr int TMEM_COLS = 2 * N < 32 ? 32 : 2 * N;
    // two CTAs per SM (two launches in flight) while 2 x (stages + TMEM columns) fit; N = 256 fills TMEM on its own
    static constexpr int CTAS = N <= 128 ? 2 : 1;
    static constexpr int STAGES = N <= 32 ? 5 : N == 64 ? 4 : N == 128 ? 3 : 4;
    // B has its own, deeper ring: its TMA load (an L2 round trip) is issued B_SLOTS - STAGES tiles before A's stage is
    // even free, so that it is off the loop  stage free -> zero -> scatter -> mma -> stage free  that sets the tile rate
    static constexpr int B_SLOTS = N == 16 ? 10 : N == 32 ? 7 : N == 64 ? 5 : N == 128 ? 3 : 4;
    static constexpr int RING_BYTES = STAGES * A_BYTES + B_SLOTS * B_SLOT;
    // instruction descriptor (UMMA::InstrDescriptor): D fp32, A/B fp16, A K-major, B MN-major, N >> 3, M >> 4
    static constexpr uint32_t IDESC = (1u << 4) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
};

template <int N> constexpr int tile_stages() { return b_geom<N>::STAGES; }

template <int N, bool DBG> __global__ void __launch_bounds__(THREADS, b_geom<N>::CTAS) k_b200_tile(tile_args a, const __grid_constant__ CUtensorMap b_map)
{
    using G = b_geom<N>;
    constexpr int S = G::STAGES, SB = G::B_SLOTS;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bars[3 * S + 4 + SB];
    __shared__ uint32_t tmem_slot;
    // 1024-byte alignment of the stage buffers (128-byte swizzle atoms)
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    unsigned char *smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
    uint32_t *s_ptr = (uint32_t *)(smem_al + (size_t)G::RING_BYTES); // the CTA's slice of tile_ptr
    const uint32_t bring0 = smem0 + S * A_BYTES;                  // B's ring behind A's stages

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t n_cta = gridDim.x, cta = blockIdx.x;
    const uint32_t t0 = (uint32_t)((uint64_t)cta * a.n_tiles / n_cta), t1 = (uint32_t)((uint64_t)(cta + 1) * a.n_tiles / n_cta);
    const uint32_t n_local = t1 - t0;
    const uint32_t KT = a.k_tiles;

    if (DBG && threadIdx.x == 0) stamp(a, 0);
    const uint32_t bar0 = smem_u32(bars);
    // full[s] = bar0 + 8 s, empty[s] = + 8 (S + s), zeroed[s] = + 8 (2 S + s), acc_full[b] = + 8 (3 S + b), acc_empty[b] = + 8 (3 S + 2 + b)
    const uint32_t full0 = bar0, empty0 = bar0 + 8u * S, zeroed0 = bar0 + 16u * S, accf0 = bar0 + 24u * S, acce0 = accf0 + 16u, bfull0 = accf0 + 32u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) mbar_init(&bars[s], PRODUCER_WARPS), mbar_init(&bars[S + s], 1), mbar_init(&bars[2 * S + s], 1);
        for (int b = 0; b < SB; b++) mbar_init(&bars[3 * S + 4 + b], 1);
        mbar_init(&bars[3 * S], 1), mbar_init(&bars[3 * S + 1], 1);
        mbar_init(&bars[3 * S + 2], 4), mbar_init(&bars[3 * S + 3], 4);
    }
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"((uint32_t)G::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (uint32_t i = threadIdx.x; i <= n_local; i += blockDim.x) s_ptr[i] = a.tile_ptr[t0 + i]; // the host made room for n_local + 1 entries
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    row::pdl_launch_dependents();
    if (DBG && threadIdx.x == 0) stamp(a, 1);

    auto run_begin = [&](uint32_t it) { return s_ptr[it]; };
    const uint32_t skip = DBG ? a.skip : 0u;

    if (warp < PRODUCER_WARPS) {
        // ------------------------------------------------------------------------------------------ producers
        const uint32_t gt = threadIdx.x;
        long long w_zeroed = 0, w_store = 0, w_ld = 0, w_sts = 0, w_fence = 0;
        struct pre {
            uint4 o, v;
        };
        auto load_tile = [&](pre &P, uint32_t it) {
            if (it >= n_local) return;
            const uint32_t e = run_begin(it) + gt * 8u;
            if (e < run_begin(it + 1) && !(skip & 16u)) P.o = ld_stream_v4(a.off + e), P.v = ld_stream_v4(a.val + e);
        };
        uint32_t s = 0, ph = 0;
        auto process = [&](const pre &P, uint32_t it) {
            const uint32_t abase = smem0 + s * A_BYTES;
            const uint32_t e0 = run_begin(it), e1 = run_begin(it + 1);
            long long tp = 0;
            if (DBG) mbar_wait_t(zeroed0 + 8u * s, ph, w_zeroed), tp = clock64();
            else mbar_wait(zeroed0 + 8u * s, ph); // the stage is free and A's image is all zero
            auto scatter = [&](const uint4 &oo, const uint4 &vv) {
                const uint32_t ow[4] = {oo.x, oo.y, oo.z, oo.w}, vw[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    sts16(abase + (ow[j] & 0xffffu), (unsigned short)(vw[j] & 0xffffu));
                    sts16(abase + (ow[j] >> 16), (unsigned short)(vw[j] >> 16));
                }
            };
            long long t1 = 0, t2 = 0, t3 = 0;
            if (DBG) {
                if (e0 + gt * 8u < e1) { // force the wait for the loaded pairs here, so that it is timed on its own
                    uint32_t dep = P.o.x ^ P.v.x ^ P.o.w ^ P.v.w;
                    asm volatile("" ::"r"(dep));
                }
                t1 = clock64();
            }
            if (e0 + gt * 8u < e1 && !(skip & 1u)) scatter(P.o, P.v);
            for (uint32_t e = e0 + PASS_PAIRS + gt * 8u; e < e1; e += PASS_PAIRS) scatter(ld_stream_v4(a.off + e), ld_stream_v4(a.val + e)); // tiles denser than 31 %
            if (DBG) t2 = clock64();
            if (!(skip & 32u) && !a.consumer_fence) fence_async_smem();
            if (DBG) t3 = clock64();
            __syncwarp();
            if (lane == 0) mbar_arrive(full0 + 8u * s);
            if (DBG && lane == 0 && (warp == 0 || warp == PRODUCER_WARPS - 1)) trace(a, warp == 0 ? 1 : 2, it);
            if (DBG) {
                w_ld += t1 - tp, w_sts += t2 - t1, w_fence += t3 - t2;
                w_store += clock64() - tp;
                if (it == 0 && threadIdx.x == 0) stamp(a, 2);
            }
            if (++s == S) s = 0, ph ^= 1u;
        };
        pre P[DEPTH];
#pragma unroll
        for (int d = 0; d < DEPTH; d++) load_tile(P[d], d);
        for (uint32_t it = 0; it < n_local; it += DEPTH) {
#pragma unroll
            for (int d = 0; d < DEPTH; d++) {
                if (it + d < n_local) {
                    process(P[d], it + d);
                    load_tile(P[d], it + d + DEPTH); // DEPTH - 1 tiles stay in flight
                }
            }
        }
        if (DBG && gt == 0) counter(a, 8, w_zeroed), counter(a, 9, w_store), counter(a, 10, w_ld), counter(a, 11, w_sts), counter(a, 5, w_fence);
    } else if (warp == MMA_WARP) {
        // ------------------------------------------------------------------------------------------ MMA issue
        if (lane == 0) {
            uint32_t it = 0, seg = 0, s = 0, ph = 0, sb = 0, phb = 0, kt0 = t0 % KT;
            long long w_full = 0, w_acce = 0;
            const uint64_t adesc0 = smem_desc(smem0, 16, 1024, 2u), bdesc0 = smem_desc(bring0, G::GROUP_BYTES, G::SBO, G::LAYOUT);
            while (it < n_local) {
                const uint32_t len = min(KT - kt0, n_local - it), ab = seg & 1u;
                if (DBG) mbar_wait_t(acce0 + 8u * ab, ((seg >> 1) & 1u) ^ 1u, w_acce);
                else mbar_wait(acce0 + 8u * ab, ((seg >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d = tmem_base + ab * N;
                for (uint32_t i = 0; i < len; i++) {
                    if (DBG) mbar_wait_t(bfull0 + 8u * sb, phb, w_acce);
                    else mbar_wait(bfull0 + 8u * sb, phb); // B's slice has landed (issued several tiles ago)
                    if (DBG) mbar_wait_t(full0 + 8u * s, ph, w_full);
                    else mbar_wait(full0 + 8u * s, ph);
                    if (a.consumer_fence) fence_async_smem(); // the producers' stores, acquired through full[], made visible to the tensor core's proxy
                    tc_fence_after();
                    if (DBG && it + i == 0) stamp(a, 3);
                    if (DBG) trace(a, 3, it + i);
                    const uint64_t ad = adesc0 + (uint64_t)(s * (A_BYTES >> 4)), bd = bdesc0 + (uint64_t)(sb * (G::B_SLOT >> 4));
#pragma unroll
                    for (int k = 0; k < TK / 16; k++)
                        if (!(skip & 4u)) tc_mma_f16(d, ad + (uint64_t)(2 * k), bd + (uint64_t)(k * (G::K16_STEP >> 4)), G::IDESC, (i > 0 || k > 0) ? 1u : 0u);
                    tc_commit(empty0 + 8u * s);
                    if (DBG) trace(a, 4, it + i);
                    if (++s == S) s = 0, ph ^= 1u;
                    if (++sb == SB) sb = 0, phb ^= 1u;
                }
                tc_commit(accf0 + 8u * ab);
                it += len, seg++, kt0 = 0;
            }
            if (DBG) stamp(a, 4), counter(a, 12, w_full), counter(a, 13, w_acce);
        }
        __syncwarp();
    } else if (warp == AHEAD_WARP) {
        // ------------------------------------------------------------------------------------------ L2 prefetch, zero fill, B loads
        // Zero fill with 32 x STS.128 per lane-row of the warp: measured 128 cycles per 16 KB (the full 128 B/clk), against
        // ~250 for one st.bulk (UMEMSETS), during which the producers' scatter stores queue behind it (scratch/ubench/smem.cu).
        auto pf = [&](uint32_t it) {
            if (it >= n_local) return;
            const uint32_t e0 = run_begin(it), bytes = (run_begin(it + 1) - e0) * 2u;
            if (bytes) l2_prefetch_bulk(a.off + e0, bytes), l2_prefetch_bulk(a.val + e0, bytes);
        };
        long long w_empty = 0, w_zero = 0;
        uint32_t s = 0, ph = 0, kt = t0 % KT, jb = 0, sb = 0;
        // B's 64 x N slice of tile jb into slot jb % SB: one box per 64-column atom; rows past K come back as zeros
        auto load_b = [&]() {
            if (jb < n_local && lane == 0) {
                mbar_expect_tx(bfull0 + 8u * sb, (uint32_t)G::BYTES);
#pragma unroll
                for (int g = 0; g < G::GROUPS; g++) tma_load_2d(bring0 + sb * G::B_SLOT + g * G::GROUP_BYTES, &b_map, (uint32_t)(g * G::ATOM_N), kt * TK, bfull0 + 8u * sb);
            }
            jb++;
            if (++kt == KT) kt = 0;
            if (++sb == SB) sb = 0;
        };
        if (lane == 0)
            for (uint32_t it = 0; it < PF_TILES; it++) pf(it);
        if (a.dep_wait) row::pdl_wait(); // the L2 prefetch of A's runs is out; B belongs to the predecessor until it has completed
        for (int i = 0; i < SB - S; i++) load_b(); // the slots no tile has used yet
        for (uint32_t it = 0; it < n_local; it++) {
            if (lane == 0) pf(it + PF_TILES);
            if (DBG) mbar_wait_t(empty0 + 8u * s, ph ^ 1u, w_empty);
            else mbar_wait(empty0 + 8u * s, ph ^ 1u); // the tensor core has finished reading the stage
            const long long tz = DBG ? clock64() : 0;
            if (!(skip & 2u)) {
                if (a.bulk_zero) {
                    if (lane == 0) smem_zero_bulk(smem0 + s * A_BYTES, A_BYTES);
                } else {
                    const uint32_t zb = smem0 + s * A_BYTES + lane * 16u;
#pragma unroll
                    for (int j = 0; j < A_BYTES / 512; j++) sts128(zb + j * 512, make_uint4(0, 0, 0, 0));
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(zeroed0 + 8u * s);
            if (DBG && lane == 0) trace(a, 0, it);
            load_b(); // tile it + SB - S: its slot was last read by the mma of tile it - S, which has completed
            if (DBG) w_zero += clock64() - tz;
            if (++s == S) s = 0, ph ^= 1u;
        }
        if (DBG && lane == 0) counter(a, 14, w_empty), counter(a, 15, w_zero);
    } else {
        // ------------------------------------------------------------------------------------------ epilogue
        const uint32_t q = warp - EPI_WARP0, rip = q * 32 + lane, et = threadIdx.x - EPI_WARP0 * 32;
        __half *C = (__half *)a.C;
        uint32_t it = 0, seg = 0;
        auto wait_for = [&](const uint32_t *flag, uint32_t v) { // one lane polls, the warp follows
            if (lane == 0)
                while (ld_acquire(flag) != v) __nanosleep(32);
            __syncwarp();
        };
        while (it < n_local) {
            const uint32_t t = t0 + it, panel = t / KT, kt0 = t % KT, len = min(KT - kt0, n_local - it), ab = seg & 1u;
            const bool partial_in = kt0 > 0;                 // someone else owns the panel: leave a partial
            const bool owner_wait = !partial_in && len < KT; // this CTA owns the panel and others hold its tail
            // what does not depend on the accumulator goes first, while the tile loop is still running: the hand-shakes
            // with the previous launch and the output row of this thread
            uint32_t orow = 0xffffffffu;
            if (partial_in) wait_for(a.slot_free + cta, a.seq - 1); // the previous launch's owner has read this CTA's slot
            else {
                wait_for(a.panel_seq + panel, a.seq - 1); // the previous launch has written these rows of C
                orow = a.out_row[(size_t)panel * TM + rip];
            }
            mbar_wait_idle(accf0 + 8u * ab, (seg >> 1) & 1u);
            tc_fence_after();
            if (DBG && et == 0 && it + len >= n_local) stamp(a, 7);
            const uint32_t taddr = tmem_base + ((q * 32u) << 16) + ab * N;
            if (partial_in) {
                float *slot = a.ws + ((size_t)cta * TM + rip) * N;
#pragma unroll 1
                for (int c = 0; c < N / 8; c++) {
                    uint32_t r[8];
                    tc_ld8(taddr + c * 8, r);
                    __stcg((uint4 *)(slot + c * 8), make_uint4(r[0], r[1], r[2], r[3]));
                    __stcg((uint4 *)(slot + c * 8) + 1, make_uint4(r[4], r[5], r[6], r[7]));
                }
                __threadfence();
                bar_sync(3, 128);
                if (et == 0) st_release(a.flags + cta, a.seq);
            } else {
                uint32_t c_last = cta; // contributors: the CTAs after this one whose first tile lies inside the panel
                if (owner_wait) {
                    const uint64_t panel_end = (uint64_t)(panel + 1) * KT;
                    while (c_last + 1 < n_cta && (uint64_t)(c_last + 1) * a.n_tiles / n_cta < panel_end) c_last++;
                    for (uint32_t c2 = cta + 1; c2 <= c_last; c2++) wait_for(a.flags + c2, a.seq);
                }
#pragma unroll 1
                for (int c = 0; c < N / 8; c++) {
                    uint32_t r[8];
                    tc_ld8(taddr + c * 8, r);
                    for (uint32_t c2 = cta + 1; c2 <= c_last; c2++) {
                        const float4 *slot = (const float4 *)(a.ws + ((size_t)c2 * TM + rip) * N + c * 8);
                        const float4 x = __ldcg(slot), y = __ldcg(slot + 1);
                        r[0] = __float_as_uint(__uint_as_float(r[0]) + x.x), r[1] = __float_as_uint(__uint_as_float(r[1]) + x.y);
                        r[2] = __float_as_uint(__uint_as_float(r[2]) + x.z), r[3] = __float_as_uint(__uint_as_float(r[3]) + x.w);
                        r[4] = __float_as_uint(__uint_as_float(r[4]) + y.x), r[5] = __float_as_uint(__uint_as_float(r[5]) + y.y);
                        r[6] = __float_as_uint(__uint_as_float(r[6]) + y.z), r[7] = __float_as_uint(__uint_as_float(r[7]) + y.w);
                    }
                    if (orow != 0xffffffffu) {
                        uint32_t h[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            __half2 p = __floats2half2_rn(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
                            h[j] = *(uint32_t *)&p;
                        }
                        *(uint4 *)(C + (size_t)orow * N + c * 8) = make_uint4(h[0], h[1], h[2], h[3]);
                    }
                }
                __threadfence();
                bar_sync(3, 128);
                if (et == 0) {
                    st_release(a.panel_seq + panel, a.seq);
                    for (uint32_t c2 = cta + 1; c2 <= c_last; c2++) st_release(a.slot_free + c2, a.seq);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acce0 + 8u * ab);
            it += len, seg++;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (DBG && threadIdx.x == 0) stamp(a, 6);
    if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G::TMEM_COLS) : "memory");
}

} // namespace tile
} // namespace gsd
