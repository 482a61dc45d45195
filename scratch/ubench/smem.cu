// micro-benchmarks of the shared-memory costs the tile kernel pays per tile (one CTA per SM, 1 block)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sts16(uint32_t addr, unsigned short v) { asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(v) : "memory"); }
__device__ __forceinline__ void zero_bulk(uint32_t addr, uint32_t bytes) { asm volatile("st.bulk.weak.shared::cta [%0], %1, 0;" ::"r"(addr), "l"((uint64_t)bytes) : "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// mode 0: scatter only (10 warps x 8 STS.U16 per iteration, random offsets)
// mode 1: scatter + fence.proxy.async per iteration
// mode 2: warp 10 lane 0 issues st.bulk 16 KB back to back (with a membar after each), others idle
// mode 3: scatter (warps 0-9) while warp 10 zeroes another buffer continuously
// mode 4: 4 warps zero 16 KB with STS.128 per iteration
__global__ void k(int mode, int iters, const uint16_t *offs, long long *out)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    const uint32_t base = smem_u32(sm);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t o[8];
    for (int j = 0; j < 8; j++) o[j] = offs[(threadIdx.x * 8 + j) % 4096] & 0x3ffe;
    __syncthreads();
    long long t0 = clock64();
    if (mode == 0 || mode == 1 || mode == 3) {
        if (warp < 10) {
            for (int i = 0; i < iters; i++) {
                const uint32_t b = base + (i & 3) * 16384;
#pragma unroll
                for (int j = 0; j < 8; j++) sts16(b + o[j], (unsigned short)(i + j));
                if (mode == 1) fence_async();
            }
        } else if (mode == 3 && warp == 10 && lane == 0) {
            for (int i = 0; i < iters; i++) {
                zero_bulk(base + 65536 + (i & 1) * 16384, 16384);
            }
            asm volatile("membar.cta;" ::: "memory");
        }
    } else if (mode == 2) {
        if (warp == 10 && lane == 0) {
            for (int i = 0; i < iters; i++) {
                zero_bulk(base + (i & 3) * 16384, 16384);
                asm volatile("membar.cta;" ::: "memory");
            }
        }
    } else if (mode == 4) {
        if (warp < 4) {
            for (int i = 0; i < iters; i++) {
                const uint32_t b = base + (i & 3) * 16384;
#pragma unroll
                for (int j = 0; j < 8; j++) asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(b + (threadIdx.x + j * 128) * 16), "r"(0) : "memory");
            }
        }
    } else if (mode == 5) { // st.bulk without a membar in between (pure issue + queue)
        if (warp == 10 && lane == 0) {
            for (int i = 0; i < iters; i++) zero_bulk(base + (i & 3) * 16384, 16384);
            asm volatile("membar.cta;" ::: "memory");
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
int main()
{
    uint16_t h[4096];
    uint32_t s = 12345;
    for (int i = 0; i < 4096; i++) { s = s * 1664525u + 1013904223u; h[i] = (uint16_t)((s >> 8) & 0x3ffe); }
    // conflict-free variant: lane l of a warp hits bank l
    uint16_t hc[4096];
    for (int i = 0; i < 4096; i++) { int t = i / 8, j = i % 8, lane = t % 32, w = t / 32; hc[i] = (uint16_t)(((w * 8 + j) * 128 + lane * 4) & 0x3ffe); }
    uint16_t *d, *dc; long long *out;
    cudaMalloc(&d, sizeof h); cudaMalloc(&dc, sizeof hc); cudaMalloc(&out, 8 * 148);
    cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); cudaMemcpy(dc, hc, sizeof hc, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const int iters = 2000;
    const char *names[] = {"scatter 80 STS.U16/iter (random banks)", "scatter + fence.proxy.async", "st.bulk 16 KB + membar", "scatter while st.bulk runs", "STS.128 zero 16 KB (4 warps)", "st.bulk 16 KB back to back"};
    for (int pass = 0; pass < 2; pass++)
        for (int mode = 0; mode < 6; mode++) {
            for (int blocks = 1; blocks <= 2; blocks++) {
                // blocks=2 on the same SM is not controllable here; run a full grid of 148 or 296 instead
                int grid = 148 * blocks;
                k<<<grid, 352, 100 * 1024>>>(mode, iters, pass ? dc : d, out);
                cudaDeviceSynchronize();
                long long r[296];
                cudaMemcpy(r, out, 8 * 148, cudaMemcpyDeviceToHost);
                printf("%s offsets, %-42s grid %3d: %8.1f cycles/iter\n", pass ? "conflict-free" : "random      ", names[mode], grid, (double)r[0] / iters);
            }
        }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
