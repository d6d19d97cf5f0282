// kernels.cu — sm_100a kernels of the DPF-PIR answer path.
//
//   top_kernel        root -> frontier of CTA tasks (levels 0..Fc), one CTA, breadth first
//   fused_kernel      per CTA task: breadth-first to one node per thread, then a per-thread
//                     depth-first walk in lockstep (sibling stack in shared memory) to the leaves;
//                     every 32 selection bits the warp XOR-accumulates the matching database rows
//                     with coalesced 128-bit loads; bits never leave the SM            (SCAN mode)
//                     or the bits are written out one byte per row like Go's []bool   (BITS mode)
//   eval_points       (*Dpf).Evaluate2P on arbitrary points (dpf API, keyword queries)
//   scan_bytes        masked row XOR from a byte-per-row bit vector (WithExpandedBits, keyword mode)
//   synth_fill        seeded synthetic database
//
// Reference semantics: dpf/server.go:55-101, dpf/common.go:56-75, db.go:75-174, slot.go:19-30.
#include "kernels.cuh"
#include <atomic>

namespace pirb {

static std::atomic<uint64_t> g_launches{0};
uint64_t launch_count() { return g_launches.load(); }
void count_launch() { g_launches++; }

extern __shared__ __align__(16) uint8_t g_smem[];

// Every table-using kernel is launched with dynamic shared memory reaching shared address
// kTableEnd; the scratch area is [smem_addr(g_smem), kTableBase).
__device__ __forceinline__ void require_layout() {
    uint32_t base = smem_addr(g_smem);
    uint32_t dyn;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
    if (base + dyn < kTableEnd || base >= kTableBase) __trap();
}

// ------------------------------------------------------------------------------------------------
// top kernel
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTopThreads, 1)
top_kernel(const __grid_constant__ QueryConst qc, const __grid_constant__ TopPlan tp,
           const uint32_t* __restrict__ t0) {
    require_layout();
    fill_tables(t0);
    const uint32_t lc = lane_const();
    const uint64_t task_last = tp.task_first + tp.task_count - 1;
    if (threadIdx.x == 0) {
        tp.buf_s[0][0] = make_uint4(qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]);
        tp.buf_t[0][0] = (uint8_t)qc.t_init;
    }
    __syncthreads();
    for (uint32_t i = 0; i < tp.Fc; i++) {
        const uint32_t sh = tp.Fc - i;
        const uint64_t lo = tp.task_first >> sh, hi = task_last >> sh;
        const uint64_t lo_n = tp.task_first >> (sh - 1), hi_n = task_last >> (sh - 1);
        const uint4* __restrict__ in_s = tp.buf_s[i & 1];
        const uint8_t* __restrict__ in_t = tp.buf_t[i & 1];
        uint4* out_s = tp.buf_s[(i + 1) & 1];
        uint8_t* out_t = tp.buf_t[(i + 1) & 1];
        for (uint64_t p = lo + threadIdx.x; p <= hi; p += blockDim.x) {
            uint4 sv = in_s[p - lo];
            uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
            uint32_t t = in_t[p - lo];
            Children ch;
            expand_node<1>(qc, s, t, i, lc, ch);
            uint64_t cl = 2 * p, cr = 2 * p + 1;
            if (cl >= lo_n && cl <= hi_n) {
                out_s[cl - lo_n] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                out_t[cl - lo_n] = (uint8_t)ch.tl;
            }
            if (cr >= lo_n && cr <= hi_n) {
                out_s[cr - lo_n] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                out_t[cr - lo_n] = (uint8_t)ch.tr;
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// eval_points: one root-to-leaf walk per thread
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
eval_points_kernel(const __grid_constant__ QueryConst qc, const uint64_t* __restrict__ xs,
                   uint64_t x_first, uint64_t n, int64_t* __restrict__ values,
                   uint8_t* __restrict__ bits, const uint32_t* __restrict__ t0) {
    require_layout();
    fill_tables(t0);
    __syncthreads();
    const uint32_t lc = lane_const();
    const uint32_t nb = qc.num_bits;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t x = xs ? xs[i] : x_first + i;
        uint32_t s[4] = {qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]};
        uint32_t t = qc.t_init;
        for (uint32_t lvl = 0; lvl < nb; lvl++) {
            Children ch;
            expand_node<1>(qc, s, t, lvl, lc, ch);
            uint32_t bit = (uint32_t)(x >> (nb - 1 - lvl)) & 1u;  // getBit, dpf/common.go:56-58
#pragma unroll
            for (int c = 0; c < 4; c++) s[c] = bit ? ch.sr[c] : ch.sl[c];
            t = bit ? ch.tr : ch.tl;
        }
        if (values) values[i] = leaf_value(s[0], s[1], t, qc.final_cw, qc.share_number);
        if (bits) bits[i] = (uint8_t)leaf_select_bit(s[0], s[1], t, qc.fcw_parity);
    }
}

// ------------------------------------------------------------------------------------------------
// synthetic database: the "pir_synth" generator of DESIGN.md (the test-suite checks it against its CPU twin)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t synth_word(uint64_t seed, uint64_t slot, uint32_t w) {
    uint64_t z = seed + slot * 0x9E3779B97F4A7C15ull + (uint64_t)w * 0xD1B54A32D192ED03ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void synth_fill_kernel(uint8_t* __restrict__ db, uint64_t slot_begin, uint64_t n_slots,
                                  uint32_t slot_bytes, uint32_t pitch, uint64_t seed) {
    // one thread per 8-byte word of the pitched layout; bytes past slot_bytes are zero padding
    const uint32_t wps = pitch / 8;
    const uint64_t total = n_slots * wps;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total;
         g += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t slot = g / wps;
        uint32_t w = (uint32_t)(g % wps);
        uint64_t v = 0;
        if (w * 8u < slot_bytes) {
            v = synth_word(seed, slot_begin + slot, w);
            uint32_t valid = slot_bytes - w * 8u;
            if (valid < 8u) v &= (1ull << (8u * valid)) - 1ull;
        }
        reinterpret_cast<uint64_t*>(db)[g] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// fused kernel
// ------------------------------------------------------------------------------------------------
#ifndef PIRB_SKIP_ZERO_ROWS
#define PIRB_SKIP_ZERO_ROWS 1  // 1: rows whose selection bit is 0 are not fetched (like db.go:93)
#endif
#ifndef PIRB_DFS_ROUND_UNROLL
#define PIRB_DFS_ROUND_UNROLL 1
#endif

// streaming 128-bit load, predicated in PTX so that unrolled loops stay branch-free
__device__ __forceinline__ uint4 ldg_stream(const uint8_t* p, bool pred) {
    uint4 v;
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
        "@p ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
        : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
        : "l"(p), "r"((uint32_t)pred));
    return v;
}

// bits k in [0, rw) of a lane stream starting at global row `rowbase` that fall inside the shard
__device__ __forceinline__ uint32_t range_mask(uint64_t rowbase, uint32_t rw, uint64_t first_row,
                                               uint64_t end_row) {
    if (rowbase >= end_row || rowbase + rw <= first_row) return 0u;
    uint32_t kmin = rowbase >= first_row ? 0u : (uint32_t)(first_row - rowbase);
    uint32_t kmax = (rowbase + rw <= end_row) ? rw : (uint32_t)(end_row - rowbase);
    uint32_t hi = kmax >= 32u ? 0xffffffffu : ((1u << kmax) - 1u);
    uint32_t lo = kmin >= 32u ? 0xffffffffu : ((1u << kmin) - 1u);
    return hi & ~lo;
}

template <int NACC, bool BITS>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_kernel(const __grid_constant__ QueryConst qc, const __grid_constant__ FusedPlan fp,
             const uint32_t* __restrict__ t0) {
    require_layout();
    fill_tables(t0);
    const uint32_t lc = lane_const();
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t D = fp.D, B = fp.B;
    const uint32_t nb = qc.num_bits;
    const uint32_t lvl0 = fp.Fc + B;            // level of the per-thread subtree roots
    const uint32_t RW = D >= 4 ? 32u : (2u << D);  // rows per bit word
    const uint32_t nlp = 1u << D;               // leaf parents per thread
    const uint64_t end_row = fp.first_row + fp.n_rows;
    const uint32_t c = (uint32_t)(fp.row_bytes >> 4);
    const uint32_t logc = 31u - __clz(c);       // used when NACC == 1 (c is a power of two <= 32)

    uint32_t* stk = reinterpret_cast<uint32_t*>(g_smem);
    uint4* bfs_s = reinterpret_cast<uint4*>(g_smem);                 // [2][512]
    uint8_t* bfs_t = g_smem + 2 * kFusedThreads * sizeof(uint4);     // [2][512]

    uint4 acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[k] = make_uint4(0, 0, 0, 0);

    for (uint64_t task = blockIdx.x; task < fp.task_count; task += gridDim.x) {
        __syncthreads();  // tables filled / previous task's stack no longer in use
        if (tid == 0) {
            bfs_s[0] = fp.frontier_s[task];
            bfs_t[0] = fp.frontier_t[task];
        }
        __syncthreads();
        for (uint32_t b = 0; b < B; b++) {
            if (tid < (1u << b)) {
                uint4 sv = bfs_s[(b & 1) * kFusedThreads + tid];
                uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
                uint32_t t = bfs_t[(b & 1) * kFusedThreads + tid];
                Children ch;
                expand_node<1>(qc, s, t, fp.Fc + b, lc, ch);
                uint32_t o = ((b + 1) & 1) * kFusedThreads + 2 * tid;
                bfs_s[o] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                bfs_s[o + 1] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                bfs_t[o] = (uint8_t)ch.tl;
                bfs_t[o + 1] = (uint8_t)ch.tr;
            }
            __syncthreads();
        }
        uint32_t s[4] = {0, 0, 0, 0}, t = 0;
        const bool active = tid < (1u << B);
        if (active) {
            uint4 sv = bfs_s[(B & 1) * kFusedThreads + tid];
            s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
            t = bfs_t[(B & 1) * kFusedThreads + tid];
        }
        __syncthreads();  // the stack aliases the BFS buffers

        // node index of this thread's subtree root at level lvl0, and its first global row
        const uint64_t f = ((fp.task_first + task) << B) + tid;
        const uint64_t f_warp0 = ((fp.task_first + task) << B) + (tid & ~31u);
        const uint32_t rows_log = nb - lvl0;  // = D + 1
        const uint64_t warp_row0 = f_warp0 << rows_log;
        const uint64_t warp_row1 = (f_warp0 + 32) << rows_log;
        if (warp * 32u >= (1u << B) && (1u << B) >= 32u) continue;          // idle warp
        if ((1u << B) < 32u && warp > 0) continue;
        if (warp_row0 >= end_row || warp_row1 <= fp.first_row) continue;    // outside the shard

        uint32_t tst = 0, d = 0, bw = 0, i = 0;
        while (true) {
            Children ch;
            expand_node<PIRB_DFS_ROUND_UNROLL>(qc, s, t, lvl0 + d, lc, ch);
            if (d < D) {
                // push the right child, descend left
#pragma unroll
                for (int w = 0; w < 4; w++) stk[(d * 4 + w) * kFusedThreads + tid] = ch.sr[w];
                tst = (tst & ~(1u << d)) | (ch.tr << d);
#pragma unroll
                for (int w = 0; w < 4; w++) s[w] = ch.sl[w];
                t = ch.tl;
                d++;
                continue;
            }
            // leaf parent: two selection bits
            uint32_t bl = leaf_select_bit(ch.sl[0], ch.sl[1], ch.tl, qc.fcw_parity);
            uint32_t br = leaf_select_bit(ch.sr[0], ch.sr[1], ch.tr, qc.fcw_parity);
            bw |= (bl | (br << 1)) << ((2u * i) & 31u);
            i++;
            if ((i & 15u) == 0u || i == nlp) {
                const uint32_t j = (i - 1u) >> 4;  // word index inside the thread's subtree
                const uint64_t rowbase = (f << rows_log) + 32ull * j;
                uint32_t rmask = active ? range_mask(rowbase, RW, fp.first_row, end_row) : 0u;
                if (BITS) {
                    for (uint32_t k = 0; k < RW; k++)
                        if ((rmask >> k) & 1u)
                            fp.bits_out[rowbase + k - fp.first_row] = (uint8_t)((bw >> k) & 1u);
                } else {
                    const uint32_t sel = bw & rmask;
                    const uint32_t ldm = PIRB_SKIP_ZERO_ROWS ? sel : rmask;
                    for (uint32_t sl = 0; sl < 32u; sl++) {
                        const uint32_t w = __shfl_sync(0xffffffffu, sel, sl);
                        const uint32_t lm = __shfl_sync(0xffffffffu, ldm, sl);
                        if (lm == 0u) continue;
                        const uint64_t rb = ((f_warp0 + sl) << rows_log) + 32ull * j;
                        const uint8_t* base = fp.db + (rb - fp.first_row) * fp.row_bytes;
                        // note: rb may be < first_row for masked-out leading rows; those are never
                        // dereferenced because their bits in lm are 0.
                        if (NACC == 1) {
                            const uint32_t rpi = 32u >> logc;
                            const uint32_t iters = ((RW << logc) + 31u) >> 5;
                            const uint32_t rsub = lane >> logc;
                            const uint8_t* p = base + (uint64_t)rsub * fp.row_bytes + ((lane & (c - 1u)) << 4);
#pragma unroll 8
                            for (uint32_t it = 0; it < iters; it++) {
                                const uint32_t r = it * rpi + rsub;
                                const bool ld = r < RW && ((lm >> r) & 1u);
                                uint4 v = ldg_stream(p + (uint64_t)it * rpi * fp.row_bytes, ld);
                                const uint32_t m = (r < RW) ? (0u - ((w >> r) & 1u)) : 0u;
                                acc[0].x ^= v.x & m; acc[0].y ^= v.y & m;
                                acc[0].z ^= v.z & m; acc[0].w ^= v.w & m;
                            }
                        } else {
                            const uint8_t* p = base + (lane << 4);
#pragma unroll 2
                            for (uint32_t r = 0; r < RW; r++) {
                                const bool ld = (lm >> r) & 1u;
                                const uint32_t m = 0u - ((w >> r) & 1u);
#pragma unroll
                                for (int k = 0; k < NACC; k++) {
                                    uint4 v = ldg_stream(p + (uint64_t)r * fp.row_bytes + k * 512, ld);
                                    acc[k].x ^= v.x & m; acc[k].y ^= v.y & m;
                                    acc[k].z ^= v.z & m; acc[k].w ^= v.w & m;
                                }
                            }
                        }
                    }
                }
                bw = 0;
            }
            if (i == nlp) break;
            // pop: next subtree root is the right sibling stacked at depth D - ctz(i)
            d = D - (uint32_t)(__ffs(i) - 1);
#pragma unroll
            for (int w = 0; w < 4; w++) s[w] = stk[((d - 1) * 4 + w) * kFusedThreads + tid];
            t = (tst >> (d - 1)) & 1u;
        }
    }

    if (!BITS) {
        // CTA reduction of the accumulators, then one atomic XOR per answer word
        __syncthreads();
        uint32_t* red = reinterpret_cast<uint32_t*>(g_smem);
        const uint32_t words = (uint32_t)(fp.row_bytes >> 2);
        for (uint32_t k = tid; k < words; k += blockDim.x) red[k] = 0;
        __syncthreads();
        if (NACC == 1) {
            for (uint32_t off = 16; off >= c && off > 0; off >>= 1) {
                acc[0].x ^= __shfl_xor_sync(0xffffffffu, acc[0].x, off);
                acc[0].y ^= __shfl_xor_sync(0xffffffffu, acc[0].y, off);
                acc[0].z ^= __shfl_xor_sync(0xffffffffu, acc[0].z, off);
                acc[0].w ^= __shfl_xor_sync(0xffffffffu, acc[0].w, off);
            }
            if (lane < c) {
                atomicXor(&red[lane * 4 + 0], acc[0].x);
                atomicXor(&red[lane * 4 + 1], acc[0].y);
                atomicXor(&red[lane * 4 + 2], acc[0].z);
                atomicXor(&red[lane * 4 + 3], acc[0].w);
            }
        } else {
#pragma unroll
            for (int k = 0; k < NACC; k++) {
                uint32_t col = 32u * k + lane;
                atomicXor(&red[col * 4 + 0], acc[k].x);
                atomicXor(&red[col * 4 + 1], acc[k].y);
                atomicXor(&red[col * 4 + 2], acc[k].z);
                atomicXor(&red[col * 4 + 3], acc[k].w);
            }
        }
        __syncthreads();
        for (uint32_t k = tid; k < words; k += blockDim.x) {
            uint32_t v = red[k];
            if (v) atomicXor(&fp.out[k], v);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// scan_bytes: out ^= XOR of rows r with bits[r] != 0        (db.go:91-104, slot.go:19-30)
// Any row size that is a multiple of 16 bytes.  blockIdx.x = column block, blockIdx.y = row group.
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256;

__global__ void __launch_bounds__(kScanThreads)
scan_bytes_kernel(const uint8_t* __restrict__ db, uint64_t row_bytes, uint64_t n_rows,
                  const uint8_t* __restrict__ bits, uint32_t* __restrict__ out, uint32_t cpb,
                  uint32_t rpi, uint64_t rows_per_block) {
    __shared__ uint32_t red[kScanThreads * 4];
    const uint32_t t = threadIdx.x;
    const uint32_t c = (uint32_t)(row_bytes >> 4);
    for (uint32_t k = t; k < cpb * 4u; k += blockDim.x) red[k] = 0;
    __syncthreads();
    const uint32_t cidx = t % cpb, rsub = t / cpb;
    const uint32_t col = blockIdx.x * cpb + cidx;
    uint4 acc = make_uint4(0, 0, 0, 0);
    if (rsub < rpi && col < c) {
        const uint64_t r0 = (uint64_t)blockIdx.y * rows_per_block;
        uint64_t r1 = r0 + rows_per_block;
        if (r1 > n_rows) r1 = n_rows;
        const uint8_t* p = db + ((uint64_t)col << 4);
#pragma unroll 4
        for (uint64_t r = r0 + rsub; r < r1; r += rpi) {
            uint4 v = ldg_stream(p + r * row_bytes, bits[r] != 0);
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
        }
        atomicXor(&red[cidx * 4 + 0], acc.x);
        atomicXor(&red[cidx * 4 + 1], acc.y);
        atomicXor(&red[cidx * 4 + 2], acc.z);
        atomicXor(&red[cidx * 4 + 3], acc.w);
    }
    __syncthreads();
    for (uint32_t k = t; k < cpb * 4u; k += blockDim.x) {
        uint32_t colw = blockIdx.x * cpb * 4u + k;
        uint32_t v = red[k];
        if (colw < c * 4u && v) atomicXor(&out[colw], v);
    }
}

__global__ void probe_smem_kernel(uint32_t* out) { *out = smem_addr(g_smem); }

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
bool fused_supports_row_bytes(uint64_t row_bytes) {
    if (row_bytes == 0 || (row_bytes & 15)) return false;
    uint64_t c = row_bytes >> 4;
    if (c & (c - 1)) return false;
    return c <= 256;
}

cudaError_t probe_smem_base(uint32_t* out_base) {
    uint32_t* d = nullptr;
    cudaError_t e = cudaMalloc(&d, 4);
    if (e != cudaSuccess) return e;
    probe_smem_kernel<<<1, 32, 1024>>>(d);
    g_launches++;
    e = cudaMemcpy(out_base, d, 4, cudaMemcpyDeviceToHost);
    cudaFree(d);
    return e;
}

template <typename K>
static cudaError_t set_max_smem(K kernel) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTableEnd);
}

cudaError_t configure_kernels() {
    cudaError_t e;
    if ((e = set_max_smem(top_kernel)) != cudaSuccess) return e;
    if ((e = set_max_smem(eval_points_kernel)) != cudaSuccess) return e;
#define PIRB_CFG(N)                                                                  \
    if ((e = set_max_smem(fused_kernel<N, false>)) != cudaSuccess) return e;         \
    if ((e = set_max_smem(fused_kernel<N, true>)) != cudaSuccess) return e;
    PIRB_CFG(1) PIRB_CFG(2) PIRB_CFG(4) PIRB_CFG(8)
#undef PIRB_CFG
    e = configure_answer_kernel();
    if (e == cudaSuccess) e = configure_batch_scan();
    if (e == cudaSuccess) e = configure_keyword_kernel();
    return e;
}

cudaError_t launch_top(const QueryConst& qc, const TopPlan& tp, const uint32_t* d_t0, uint32_t smem_base,
                       cudaStream_t st) {
    top_kernel<<<1, kTopThreads, kTableEnd - smem_base, st>>>(qc, tp, d_t0);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_fused(const QueryConst& qc, const FusedPlan& fp, bool bits_mode, const uint32_t* d_t0,
                         uint32_t smem_base, int num_sms, cudaStream_t st) {
    uint64_t c = fp.row_bytes >> 4;
    int nacc = c <= 32 ? 1 : (int)(c / 32);
    unsigned grid = (unsigned)(fp.task_count < (uint64_t)num_sms ? fp.task_count : (uint64_t)num_sms);
    if (grid == 0) return cudaSuccess;
    size_t smem = kTableEnd - smem_base;
#define PIRB_LAUNCH(N)                                                                             \
    if (bits_mode)                                                                                 \
        fused_kernel<N, true><<<grid, kFusedThreads, smem, st>>>(qc, fp, d_t0);                    \
    else                                                                                           \
        fused_kernel<N, false><<<grid, kFusedThreads, smem, st>>>(qc, fp, d_t0);
    switch (nacc) {
        case 1: PIRB_LAUNCH(1) break;
        case 2: PIRB_LAUNCH(2) break;
        case 4: PIRB_LAUNCH(4) break;
        case 8: PIRB_LAUNCH(8) break;
        default: return cudaErrorInvalidValue;
    }
#undef PIRB_LAUNCH
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_eval_points(const QueryConst& qc, const uint64_t* xs, uint64_t x_first, uint64_t n,
                               int64_t* values, uint8_t* bits, const uint32_t* d_t0, uint32_t smem_base,
                               cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint64_t blocks = (n + 1023) / 1024;
    if (blocks > (uint64_t)sms) blocks = sms;
    eval_points_kernel<<<(unsigned)blocks, 1024, kTableEnd - smem_base, st>>>(qc, xs, x_first, n, values,
                                                                             bits, d_t0);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_scan_bytes(const uint8_t* db, uint64_t row_bytes, uint64_t n_rows, const uint8_t* bits,
                              uint32_t* out, int num_sms, cudaStream_t st) {
    if (n_rows == 0) return cudaSuccess;
    uint64_t c = row_bytes >> 4;
    uint32_t cpb = (uint32_t)(c < (uint64_t)kScanThreads ? c : (uint64_t)kScanThreads);
    uint32_t rpi = kScanThreads / cpb;
    uint32_t col_blocks = (uint32_t)((c + cpb - 1) / cpb);
    // enough row groups to fill the machine a few times over, at least rpi*8 rows each
    uint64_t want = (uint64_t)num_sms * 8 / col_blocks + 1;
    uint64_t rows_per_block = (n_rows + want - 1) / want;
    uint64_t min_rows = (uint64_t)rpi * 8;
    if (rows_per_block < min_rows) rows_per_block = min_rows;
    uint64_t row_groups = (n_rows + rows_per_block - 1) / rows_per_block;
    if (row_groups > 65535) {
        row_groups = 65535;
        rows_per_block = (n_rows + row_groups - 1) / row_groups;
    }
    dim3 grid(col_blocks, (unsigned)row_groups);
    scan_bytes_kernel<<<grid, kScanThreads, 0, st>>>(db, row_bytes, n_rows, bits, out, cpb, rpi,
                                                     rows_per_block);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_synth_fill(uint8_t* db, uint64_t slot_begin, uint64_t n_slots, uint32_t slot_bytes,
                              uint32_t pitch, uint64_t seed, cudaStream_t st) {
    if (n_slots == 0) return cudaSuccess;
    uint64_t total = n_slots * (pitch / 8);
    uint64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    synth_fill_kernel<<<(unsigned)blocks, 256, 0, st>>>(db, slot_begin, n_slots, slot_bytes, pitch, seed);
    g_launches++;
    return cudaGetLastError();
}

// ---- packed selection bits -> one byte per row (Go's []bool, db.go:110-174) ------------------------
__global__ void unpack_bits_kernel(const uint32_t* __restrict__ packed, uint8_t* __restrict__ bytes, uint64_t n_rows) {
    // one thread per 4 rows: a 32-bit store of four 0/1 bytes
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g * 4 < n_rows; g += stride) {
        const uint32_t w = packed[g >> 3] >> ((g & 7) * 4);
        const uint32_t v = (w & 1u) | ((w & 2u) << 7) | ((w & 4u) << 14) | ((w & 8u) << 21);
        if (g * 4 + 4 <= n_rows) {
            *reinterpret_cast<uint32_t*>(bytes + g * 4) = v;
        } else {
            for (uint64_t k = 0; g * 4 + k < n_rows; k++) bytes[g * 4 + k] = (uint8_t)((v >> (8 * k)) & 1u);
        }
    }
}

cudaError_t launch_unpack_bits(const uint32_t* packed, uint8_t* bytes, uint64_t n_rows, cudaStream_t st) {
    if (n_rows == 0) return cudaSuccess;
    uint64_t blocks = ((n_rows + 3) / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    unpack_bits_kernel<<<(unsigned)blocks, 256, 0, st>>>(packed, bytes, n_rows);
    g_launches++;
    return cudaGetLastError();
}

// ---- multi-GPU combine signalling --------------------------------------------------
// A shard's answer kernel XORs its partial straight into the combine slot (possibly a peer GPU's memory,
// red.global.xor over NVLink).  signal_arrival runs after it on the same stream: the partial is complete,
// so it bumps the slot's arrival counter with system scope.  wait_arrivals runs on the OWNER GPU only and
// spins (bounded) until all shards have arrived; the waiting kernel and the signalling kernels run on
// different GPUs, so none of them can starve the other.
__global__ void signal_arrival_kernel(uint32_t* arrivals) {
    __threadfence_system();
    atomicAdd_system(arrivals, 1u);
}

__global__ void wait_arrivals_kernel(const uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles,
                                     uint32_t* timed_out) {
    const long long t0 = clock64();
    while (true) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(arrivals) : "memory");
        if (v >= expected) break;
        if ((uint64_t)(clock64() - t0) > timeout_cycles) {
            *timed_out = 1u;
            break;
        }
        __nanosleep(200);
    }
}

// Owner GPU, one launch per query of the end-to-end path: wait (bounded) for the shards' arrivals, move the combined
// answer to the caller's mapped pinned buffer, and leave slot and counter zeroed for the slot's next query
// (atomicExch reads and clears in one go).  *h_flag = 1 on a timeout (nothing is moved then).
__global__ void wait_read_reset_kernel(uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles, uint32_t* slot,
                                       uint32_t* h_out, uint32_t* h_flag, uint32_t words) {
    __shared__ uint32_t ok;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        ok = 1u;
        while (true) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(arrivals) : "memory");
            if (v >= expected) break;
            if ((uint64_t)(clock64() - t0) > timeout_cycles) {
                ok = 0u;
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    if (ok) {
        __threadfence_system();
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) h_out[i] = atomicExch(&slot[i], 0u);
        if (threadIdx.x == 0) *arrivals = 0u;
    }
    if (threadIdx.x == 0) *h_flag = ok ? 0u : 1u;
    __threadfence_system();
}

cudaError_t launch_wait_read_reset(uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles, uint32_t* slot, uint32_t* h_out,
                                   uint32_t* h_flag, uint32_t words, cudaStream_t st) {
    wait_read_reset_kernel<<<1, 128, 0, st>>>(arrivals, expected, timeout_cycles, slot, h_out, h_flag, words);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_signal_arrival(uint32_t* arrivals, cudaStream_t st) {
    signal_arrival_kernel<<<1, 1, 0, st>>>(arrivals);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_wait_arrivals(const uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles,
                                 uint32_t* timed_out, cudaStream_t st) {
    wait_arrivals_kernel<<<1, 1, 0, st>>>(arrivals, expected, timeout_cycles, timed_out);
    g_launches++;
    return cudaGetLastError();
}

}  // namespace pirb
