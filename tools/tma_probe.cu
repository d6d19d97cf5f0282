// tma_probe.cu — development probe (not part of the product): can the row XOR of the scan be moved off the
// LSU pipe?  16 warps run conflict-free shared-memory table lookups at the rate of the AES expansion while
// other warps fetch random 256-byte rows with cp.async.bulk into shared memory and fold them into one 256-byte
// accumulator with cp.reduce.async.bulk (xor) — all in the async proxy.  Prints the lookup warps' time with and
// without the bulk traffic and the bulk rows per cycle.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_probe.bin tools/tma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kLdsWarps = 16;
constexpr int kRowBytes = 256;
constexpr int kBatchRows = 32;
constexpr int kBatchBytes = kBatchRows * kRowBytes;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t a, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_xor(uint32_t dst, uint32_t src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.reduce.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes.xor.b32 [%0], [%1], %2, [%3];" ::"r"(dst), "r"(src), "r"(bytes), "r"(bar) : "memory");
}

extern __shared__ __align__(128) uint8_t smem[];

// mode bit0: bulk loads, bit1: bulk xor reduce, bit2: lookup warps active
__global__ void __cluster_dims__(1, 1, 1) __launch_bounds__(1024, 1) probe(const uint8_t* db, uint32_t row_mask, int mode, int tma_warps, int nbatch_ring,
                                                  int iters, int tma_batches, unsigned long long* res, uint32_t* sink) {
    uint32_t* table = reinterpret_cast<uint32_t*>(smem);                  // 32 KB: [256][32] lane-replicated
    uint8_t* acc = smem + 32768;                                          // 256 B accumulator (+pad to 1 KB)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 32768 + 1024);    // per tma warp: ring x {load, red}
    uint8_t* staging = smem + 32768 + 4096;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (uint32_t i = tid; i < 8192; i += blockDim.x) table[i] = i * 2654435761u;
    if (tid < 64) reinterpret_cast<uint32_t*>(acc)[tid] = 0;
    if (tid == 0)
        for (int k = 0; k < 4 * 16 * 2; k++) mbar_init(smem_u32(&bars[k]), 1);
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncthreads();
    const long long t0 = clock64();
    if (warp < kLdsWarps) {
        if (mode & 4) {
            uint32_t a[16];
#pragma unroll
            for (int j = 0; j < 16; j++) a[j] = tid * 977u + j * 131u;
            const uint32_t base = smem_u32(table) + lane * 4u;
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    uint32_t addr = base + ((a[j] >> 8) & 0xffu) * 128u, v;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
                    a[j] = v + a[j] * 5u;
                }
            }
            uint32_t x = 0;
#pragma unroll
            for (int j = 0; j < 16; j++) x ^= a[j];
            if (x == 0x12345u) sink[0] = x;
        }
        const long long t1 = clock64();
        if (tid == 0) res[blockIdx.x * 4 + 0] = (unsigned long long)(t1 - t0);
    } else if ((int)(warp - kLdsWarps) < tma_warps && (mode & 1)) {
        const uint32_t tw = warp - kLdsWarps;
        uint8_t* st = staging + (size_t)tw * nbatch_ring * kBatchBytes;
        uint64_t* lb = bars + tw * 32;       // load barriers [ring]
        uint64_t* rb = lb + 16;              // reduce barriers [ring]
        uint32_t h = blockIdx.x * 7919u + tw * 104729u + lane * 31u;
        for (int b = 0; b < tma_batches + nbatch_ring; b++) {
            const int slot = b % nbatch_ring, round = b / nbatch_ring;
            if (b >= nbatch_ring) {
                // batch b - ring is resident in this slot: its loads must have landed; fold it
                mbar_wait(smem_u32(&lb[slot]), (round - 1) & 1);
                if (mode & 2) {
                    if (lane == 0) mbar_expect(smem_u32(&rb[slot]), kBatchBytes);
                    __syncwarp();
                    bulk_xor(smem_u32(acc), smem_u32(st + slot * kBatchBytes + lane * kRowBytes), kRowBytes, smem_u32(&rb[slot]));
                    mbar_wait(smem_u32(&rb[slot]), (round - 1) & 1);
                }
            }
            if (b < tma_batches) {
                if (lane == 0) mbar_expect(smem_u32(&lb[slot]), kBatchBytes);
                __syncwarp();
                h = h * 1664525u + 1013904223u;
                const uint32_t row = (h >> 4) & row_mask;
                bulk_load(smem_u32(st + slot * kBatchBytes + lane * kRowBytes), db + (size_t)row * kRowBytes, kRowBytes, smem_u32(&lb[slot]));
            }
        }
        const long long t1 = clock64();
        if (lane == 0) res[blockIdx.x * 4 + 1 + (tw ? 1 : 0)] = (unsigned long long)(t1 - t0);
    }
    __syncthreads();
    if (tid < 64) atomicXor(&sink[1 + tid], reinterpret_cast<uint32_t*>(acc)[tid]);
}

int main() {
    const size_t rows = 1u << 22;
    uint8_t* db;
    cudaMalloc(&db, rows * kRowBytes);
    cudaMemset(db, 0x5a, rows * kRowBytes);
    unsigned long long* res;
    uint32_t* sink;
    cudaMalloc(&res, 148 * 4 * 8);
    cudaMalloc(&sink, 4096);
    cudaMemset(sink, 0, 4096);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int iters = 4000;       // 16 warps x 4000 x 16 lookups = 1.02M wavefronts per SM
    struct Cfg { int mode, tw, ring, batches; const char* name; } cfgs[] = {
        {4, 0, 1, 0, "lookups only"},
        {1, 1, 4, 450, "bulk loads only, 1 warp, ring 4"},
        {1, 2, 4, 225, "bulk loads only, 2 warps, ring 4"},
        {1, 4, 4, 112, "bulk loads only, 4 warps, ring 4"},
        {5, 4, 4, 112, "lookups + bulk loads only, 4 warps, ring 4"},
        {3, 1, 4, 450, "bulk loads+xor, 1 warp, ring 4"},
        {3, 2, 4, 225, "bulk loads+xor, 2 warps, ring 4"},
        {3, 4, 4, 112, "bulk loads+xor, 4 warps, ring 4"},
        {5, 1, 4, 450, "lookups + bulk loads, 1 warp"},
        {7, 1, 4, 450, "lookups + bulk loads+xor, 1 warp ring 4"},
        {7, 2, 4, 225, "lookups + bulk loads+xor, 2 warps ring 4"},
        {7, 4, 4, 112, "lookups + bulk loads+xor, 4 warps ring 4"},
        {7, 4, 2, 112, "lookups + bulk loads+xor, 4 warps ring 2"},
    };
    for (auto& c : cfgs) {
        cudaMemset(res, 0, 148 * 4 * 8);
        const size_t smem_bytes = 32768 + 4096 + (size_t)(c.tw ? c.tw : 1) * c.ring * kBatchBytes;
        for (int rep = 0; rep < 2; rep++) {
            probe<<<148, 1024, smem_bytes>>>(db, (uint32_t)rows - 1, c.mode, c.tw, c.ring, iters, c.batches, res, sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: %s\n", c.name, cudaGetErrorString(e)); return 1; }
        }
        unsigned long long h[148 * 4];
        cudaMemcpy(h, res, sizeof h, cudaMemcpyDeviceToHost);
        double l = 0, t = 0;
        for (int b = 0; b < 148; b++) { l += h[b * 4]; t += h[b * 4 + 1]; }
        l /= 148; t /= 148;
        const double rows_total = (double)c.tw * c.batches * kBatchRows;
        printf("%-45s lookup warps %9.0f cyc (%.3f wavefronts/cyc)   bulk warp0 %9.0f cyc, %6.0f rows/SM -> %.1f cyc/row\n", c.name, l,
               (c.mode & 4) ? 16.0 * iters * 16 / l : 0.0, t, rows_total, rows_total > 0 ? t / rows_total : 0.0);
    }
    return 0;
}
