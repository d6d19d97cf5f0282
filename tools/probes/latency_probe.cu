// Latency probe (development aid): dependent chains of LDS / SHFL / PRMT+LDS+LOP3 by ONE warp, as in the
// single-node levels of a task top.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o latency_probe latency_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
extern __shared__ __align__(16) uint8_t smem[];
__device__ __forceinline__ uint32_t lds(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__global__ void probe(long long* out, int iters, uint32_t region) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem) + region;
    // table: entry x (256 entries) at base + x*256 + lane*4 holds a pseudo-random byte pattern
    for (uint32_t i = threadIdx.x; i < 256u * 64u; i += blockDim.x) {
        uint32_t x = i >> 6;
        uint32_t v = (x * 2654435761u) ^ (x << 13) ^ 0x9e3779b9u;
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + i * 4u), "r"(v));
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    const uint32_t lc = (lane * 4u) | ((lane * 4u + 128u) << 8) | ((base >> 16) << 16);
    uint32_t a = lane * 77u + 1u;
    long long t0, t1;
    // (1) dependent PRMT + LDS
    t0 = clock64();
    for (int i = 0; i < iters; i++) a = lds(base + (a & 0xffu) * 256u + lane * 4u);
    t1 = clock64();
    if (lane == 0) out[0] = t1 - t0;
    // (2) dependent SHFL.IDX
    t0 = clock64();
    for (int i = 0; i < iters; i++) a = __shfl_sync(0xffffffffu, a, (lane + 1u) & 31u) + 1u;
    t1 = clock64();
    if (lane == 0) out[1] = t1 - t0;
    // (3) dependent LOP3
    t0 = clock64();
    for (int i = 0; i < iters; i++) a = (a ^ (a >> 3)) + 0x1234567u;
    t1 = clock64();
    if (lane == 0) out[2] = t1 - t0;
    // (4) the quad round: 3 SHFL + 4 PRMT + 4 LDS + 2 LOP3
    const uint32_t c = lane & 3u, g0 = lane & ~3u;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
        uint32_t a1 = __shfl_sync(0xffffffffu, a, g0 + ((c + 1u) & 3u));
        uint32_t a2 = __shfl_sync(0xffffffffu, a, g0 + ((c + 2u) & 3u));
        uint32_t a3 = __shfl_sync(0xffffffffu, a, g0 + ((c + 3u) & 3u));
        uint32_t v0 = lds(base + ((a & 0xffu) << 8) + lane * 4u);
        uint32_t v1 = lds(base + (((a1 >> 8) & 0xffu) << 8) + lane * 4u + 128u);
        uint32_t v2 = lds(base + (((a2 >> 16) & 0xffu) << 8) + lane * 4u);
        uint32_t v3 = lds(base + ((a3 >> 24) << 8) + lane * 4u + 128u);
        a = v0 ^ v1 ^ v2 ^ v3 ^ 0x55aa55aau;
    }
    t1 = clock64();
    if (lane == 0) out[3] = t1 - t0;
    // (5) single-thread round: 16 lookups
    uint32_t s[4] = {a, a * 3u, a * 5u, a * 7u};
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
        uint32_t n[4];
#pragma unroll
        for (int cc = 0; cc < 4; cc++) {
            uint32_t v0 = lds(base + ((s[cc] & 0xffu) << 8) + lane * 4u);
            uint32_t v1 = lds(base + (((s[(cc + 1) & 3] >> 8) & 0xffu) << 8) + lane * 4u + 128u);
            uint32_t v2 = lds(base + (((s[(cc + 2) & 3] >> 16) & 0xffu) << 8) + lane * 4u);
            uint32_t v3 = lds(base + ((s[(cc + 3) & 3] >> 24) << 8) + lane * 4u + 128u);
            n[cc] = v0 ^ v1 ^ v2 ^ v3 ^ 0x55aa55aau;
        }
#pragma unroll
        for (int cc = 0; cc < 4; cc++) s[cc] = n[cc];
    }
    t1 = clock64();
    if (lane == 0) out[4] = t1 - t0;
    if (lane == 0) out[5] = a ^ s[0] ^ s[1] ^ s[2] ^ s[3] ^ lc;
}

// quad round by warp 0 while warps 1..15 wait at a named barrier and warps 16..19 poll an mbarrier (like the scan
// warps of answer_kernel during the first task top).  mode bit 0: pollers present, bit 1: barrier waiters present
__global__ void probe2(long long* out, int iters, uint32_t region, int mode) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem) + region;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(smem) + region + 65536u;
    for (uint32_t i = threadIdx.x; i < 256u * 64u; i += blockDim.x) {
        uint32_t x = i >> 6;
        uint32_t v = (x * 2654435761u) ^ (x << 13) ^ 0x9e3779b9u;
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + i * 4u), "r"(v));
    }
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1) : "memory");
    __syncthreads();
    if (warp == 0) {
        uint32_t a = lane * 77u + 1u;
        const uint32_t c = lane & 3u, g0 = lane & ~3u;
        long long t0 = clock64();
#pragma unroll 1
        for (int i = 0; i < iters; i++) {
            uint32_t a1 = __shfl_sync(0xffffffffu, a, g0 + ((c + 1u) & 3u));
            uint32_t a2 = __shfl_sync(0xffffffffu, a, g0 + ((c + 2u) & 3u));
            uint32_t a3 = __shfl_sync(0xffffffffu, a, g0 + ((c + 3u) & 3u));
            uint32_t v0 = lds(base + ((a & 0xffu) << 8) + lane * 4u);
            uint32_t v1 = lds(base + (((a1 >> 8) & 0xffu) << 8) + lane * 4u + 128u);
            uint32_t v2 = lds(base + (((a2 >> 16) & 0xffu) << 8) + lane * 4u);
            uint32_t v3 = lds(base + ((a3 >> 24) << 8) + lane * 4u + 128u);
            a = v0 ^ v1 ^ v2 ^ v3 ^ 0x55aa55aau;
        }
        long long t1 = clock64();
        if (lane == 0) { out[0] = t1 - t0; out[1] = a; }
        if (lane == 0) asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
        if (mode & 2) asm volatile("bar.sync 1, 512;" ::: "memory");
    } else if (warp < 16) {
        if (mode & 2) asm volatile("bar.sync 1, 512;" ::: "memory");
    } else if (mode & 1) {
        asm volatile(
            "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(bar), "r"(0) : "memory");
    }
}
int main() {
    long long* d;
    cudaMalloc(&d, 64);
    const int iters = 1000;
    for (uint32_t region : {0u, 65536u, 131072u}) {
        const int smem_bytes = region + 65536;
        cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        for (int threads : {32, 640}) {
            probe<<<1, threads, smem_bytes>>>(d, iters, region);
            long long h[8];
            cudaError_t e = cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
            printf("region %6u threads %3d (%s): LDS chain %.1f  SHFL chain %.1f  ALU chain(2 ops) %.1f  quad round %.1f  single-thread round %.1f cycles/iter\n",
                   region, threads, cudaGetErrorString(e), h[0] / (double)iters, h[1] / (double)iters, h[2] / (double)iters,
                   h[3] / (double)iters, h[4] / (double)iters);
        }
    }
    {
        const int smem_bytes = 65536 + 65536 + 64;
        cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        for (int mode = 0; mode < 4; mode++) {
            probe2<<<1, 640, smem_bytes>>>(d, iters, 65536u, mode);
            long long h[8];
            cudaError_t e = cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
            printf("quad round with %s%s(%s): %.1f cycles/iter\n", (mode & 1) ? "4 warps polling an mbarrier " : "", (mode & 2) ? "15 warps at bar.sync " : "",
                   cudaGetErrorString(e), h[0] / (double)iters);
        }
    }
    return 0;
}
