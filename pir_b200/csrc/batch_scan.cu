// batch_scan.cu — second phase of a batched answer (BASELINE configs[3]): Q queries,
// ONE pass over the database.  Phase 1 (answer_kernel<0>, BITS mode) has expanded every query's DPF into a
// packed bit vector; this kernel streams each row from HBM once and XORs it into the accumulator of every
// query that selected it:  out[q] ^= XOR_{r : bits[q][r]} row_r   (db.go:91-104 for Q bit vectors at once).
//
// One persistent CTA per SM.  A producer warp moves blocks of rows HBM -> shared memory with
// cp.async.bulk (one bulk copy per stage, mbarrier complete_tx), 4 stages in flight; here staging through
// shared memory is right because every staged byte is read Q/2 times.  512 consumer threads hold the
// accumulators in registers: thread t serves query t / tpq and the 16-byte chunks {j*tpq + t % tpq} of the
// row, so the lanes of one query read 16*tpq contiguous bytes per shared load (conflict-free).
// Bound: shared-memory -> register traffic, Q/2 * row_bytes per row (the LSU data pipe, 128 B per cycle).
#include "kernels.cuh"

namespace pirb {

extern __shared__ __align__(128) uint8_t g_bsmem[];

constexpr int kBsConsumers = 512;
constexpr int kBsThreads = kBsConsumers + 32;
constexpr uint32_t kBsStages = 4;
constexpr uint32_t kBsStageBytes = 32768;
constexpr uint32_t kBsMaxSeg = 8;  // 16-byte chunks per thread

__device__ __forceinline__ void bs_mbar_init(uint32_t a, uint32_t c) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(c) : "memory");
}
__device__ __forceinline__ void bs_mbar_arrive(uint32_t a) {
    asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void bs_mbar_expect(uint32_t a, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.release.cta.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bs_mbar_wait(uint32_t a, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(a),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bs_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}

__global__ void __launch_bounds__(kBsThreads, 1) batch_scan_kernel(const __grid_constant__ BatchScanPlan bp) {
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t sbase = smem_addr(g_bsmem);
    const uint32_t bars = sbase + kBsStages * kBsStageBytes;  // full[kBsStages], empty[kBsStages]
    if (tid == 0) {
        for (uint32_t k = 0; k < kBsStages; k++) {
            bs_mbar_init(bars + 8u * k, 1);
            bs_mbar_init(bars + 8u * (kBsStages + k), kBsConsumers / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // 32-row blocks (one selection word per query) dealt contiguously to the CTAs
    const uint64_t nblocks = (bp.n_rows + 31ull) >> 5;
    const uint64_t b0 = (uint64_t)blockIdx.x * nblocks / gridDim.x, b1 = (uint64_t)(blockIdx.x + 1) * nblocks / gridDim.x;
    const uint32_t sr = bp.stage_rows, spb = 32u / sr;  // rows per stage, stages per block
    if (tid >= kBsConsumers) {
        // ---- producer: one elected lane issues the bulk copies ----
        if (lane == 0) {
            uint32_t it = 0;
            for (uint64_t blk = b0; blk < b1; blk++) {
                for (uint32_t sub = 0; sub < spb; sub++, it++) {
                    const uint32_t s = it % kBsStages, par = (it / kBsStages) & 1u;
                    const uint64_t row0 = (blk << 5) + (uint64_t)sub * sr;
                    if (row0 >= bp.n_rows) continue;  // `it` still advances: consumers count the same way
                    const uint64_t rows = bp.n_rows - row0 < sr ? bp.n_rows - row0 : sr;
                    bs_mbar_wait(bars + 8u * (kBsStages + s), par ^ 1u);
                    bs_mbar_expect(bars + 8u * s, (uint32_t)(rows * bp.row_bytes));
                    bs_bulk_load(sbase + s * kBsStageBytes, bp.db + row0 * bp.row_bytes, (uint32_t)(rows * bp.row_bytes), bars + 8u * s);
                }
            }
        }
        return;
    }
    // ---- consumers ----
    // thread -> (row class, query, segment): when the launch has fewer queries than the CTA has room for,
    // the spare thread groups take the rows r with r % rsplit == their class, so all 512 threads stay busy
    const uint32_t seg = tid % bp.tpq, qslot = tid / bp.tpq;
    const uint32_t qper = (uint32_t)kBsConsumers / bp.tpq / bp.rsplit;  // query slots per row class
    const uint32_t q = qslot % qper, rclass = qslot / qper;
    const bool live = q < bp.n_queries;
    const uint32_t rpat = bp.rsplit >= 32u ? (1u << rclass) : ((0xffffffffu / ((1u << bp.rsplit) - 1u)) << rclass);
    const uint32_t c = (uint32_t)(bp.row_bytes >> 4);
    uint4 acc[kBsMaxSeg];
#pragma unroll
    for (int j = 0; j < (int)kBsMaxSeg; j++) acc[j] = make_uint4(0, 0, 0, 0);
    const uint32_t* qbits = bp.bits + (uint64_t)q * bp.wpq;
    uint32_t it = 0;
    for (uint64_t blk = b0; blk < b1; blk++) {
        const uint32_t w = live ? (qbits[blk] & rpat) : 0u;
        for (uint32_t sub = 0; sub < spb; sub++, it++) {
            const uint64_t row0 = (blk << 5) + (uint64_t)sub * sr;
            if (row0 >= bp.n_rows) continue;
            const uint32_t s = it % kBsStages, par = (it / kBsStages) & 1u;
            uint32_t ws = (sr >= 32u) ? w : ((w >> (sub * sr)) & ((1u << sr) - 1u));
            bs_mbar_wait(bars + 8u * s, par);
            const uint32_t stage = sbase + s * kBsStageBytes + (seg << 4);
            while (ws) {
                const uint32_t r = (uint32_t)__ffs(ws) - 1u;
                ws &= ws - 1u;
                const uint32_t p = stage + r * (uint32_t)bp.row_bytes;
#pragma unroll
                for (int j = 0; j < (int)kBsMaxSeg; j++) {
                    if (j < (int)bp.seg_chunks && (uint32_t)j * bp.tpq + seg < c) {
                        uint4 v;
                        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(p + (((uint32_t)j * bp.tpq) << 4)));
                        acc[j].x ^= v.x; acc[j].y ^= v.y; acc[j].z ^= v.z; acc[j].w ^= v.w;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) bs_mbar_arrive(bars + 8u * (kBsStages + s));
        }
    }
    if (live) {
        uint32_t* o = bp.out + (uint64_t)q * (bp.row_bytes >> 2);
#pragma unroll
        for (int j = 0; j < (int)kBsMaxSeg; j++) {
            const uint32_t ch = (uint32_t)j * bp.tpq + seg;
            if (j < (int)bp.seg_chunks && ch < c) {
                if (acc[j].x) atomicXor(&o[ch * 4 + 0], acc[j].x);
                if (acc[j].y) atomicXor(&o[ch * 4 + 1], acc[j].y);
                if (acc[j].z) atomicXor(&o[ch * 4 + 2], acc[j].z);
                if (acc[j].w) atomicXor(&o[ch * 4 + 3], acc[j].w);
            }
        }
    }
}

// Geometry for a row size: threads per query, chunks per thread, rows per stage, queries per launch.
bool batch_scan_geometry(uint64_t row_bytes, uint32_t* tpq, uint32_t* seg_chunks, uint32_t* stage_rows, uint32_t* queries_per_launch) {
    if (row_bytes < 16 || (row_bytes & 15) || row_bytes > kBsStageBytes) return false;
    const uint32_t c = (uint32_t)(row_bytes >> 4);
    uint32_t t = 1;
    while (t * kBsMaxSeg < c) t <<= 1;
    if (t > (uint32_t)kBsConsumers) return false;
    uint32_t sr = 32;
    while ((uint64_t)sr * row_bytes > kBsStageBytes) sr >>= 1;
    *tpq = t;
    *seg_chunks = (c + t - 1) / t;
    *stage_rows = sr;
    *queries_per_launch = (uint32_t)kBsConsumers / t;
    return true;
}

cudaError_t configure_batch_scan() {
    return cudaFuncSetAttribute(batch_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(kBsStages * kBsStageBytes + 2 * kBsStages * 8 + 128));
}

cudaError_t launch_batch_scan(const BatchScanPlan& bp, int num_sms, cudaStream_t st) {
    if (bp.n_rows == 0 || bp.n_queries == 0) return cudaSuccess;
    const uint64_t nblocks = (bp.n_rows + 31ull) >> 5;
    const unsigned grid = (unsigned)(nblocks < (uint64_t)num_sms ? nblocks : (uint64_t)num_sms);
    if (bp.rsplit == 0 || (bp.rsplit & (bp.rsplit - 1)) || bp.rsplit > 32 ||
        (uint64_t)bp.n_queries * bp.tpq * bp.rsplit > (uint64_t)kBsConsumers)
        return cudaErrorInvalidValue;
    batch_scan_kernel<<<grid, kBsThreads, kBsStages * kBsStageBytes + 2 * kBsStages * 8 + 128, st>>>(bp);
    count_launch();
    return cudaGetLastError();
}

}  // namespace pirb
