// answer_kernel.cu — the dominant kernel: full-domain DPF evaluation fused with the masked row XOR
// (db.go:68-174 + dpf/server.go:55-101 + slot.go:19-30), warp-specialised for sm_100a.
//
// One persistent CTA per SM, 16 "tree" warps + kScanWarps "scan" warps:
//   tree warps   per CTA task (a subtree of 2^(B+D+1) rows): thread 0 walks from the last task's
//                common ancestor to the task root, the CTA expands B levels breadth-first, then every
//                thread walks its own 2^(D+1)-row subtree depth-first (sibling stack in shared
//                memory) with the unrolled T-table AES of aes_unrolled.cuh.  The thread's selection
//                bits (one 32-bit word) go to a double-buffered word array in shared memory.
//                Tree warps never touch global memory in the steady state.
//   scan warps   turn the words into a compact list of selected rows (popc prefix sum per 32 words)
//                and stream exactly those rows from HBM with coalesced 128-bit loads (32/c rows per
//                warp load for rows of c 16-byte chunks), XOR-accumulating in registers; rows whose
//                bit is 0 are never fetched (like db.go:93).  Loads go straight to registers: the
//                shared-memory pipe is the kernel's bottleneck (AES table lookups), so staging rows
//                through shared memory (TMA / cp.async) would cost more than it hides.
// Hand-off uses four named barriers (full/empty per buffer).  The answer is reduced per CTA and
// XORed into global memory with one atomic per word.
#include "kernels.cuh"
#include "aes_unrolled.cuh"

namespace pirb {

extern __shared__ __align__(16) uint8_t g_smem2[];

constexpr int kTreeThreads = 512;
constexpr int kScanWarps = 4;
constexpr int kAnswerThreads = kTreeThreads + 32 * kScanWarps;
constexpr int kV2MaxDepth = 4;                 // one 32-bit word of selection bits per thread
constexpr uint32_t kListEntries = 1024;        // rows selected out of 32 words, per scan warp
constexpr uint32_t kListBytes = kScanWarps * kListEntries * 2;

// scratch layout below the tables (offsets from the dynamic shared memory base)
constexpr uint32_t kOffStack = 0;                                   // [D<=4][4][512] u32 = 32 KB (aliases BFS)
constexpr uint32_t kOffBfsS = 0;                                    // [2][512] uint4    = 16 KB
constexpr uint32_t kOffBfsT = 16384;                                // [2][512] u8       =  1 KB
constexpr uint32_t kOffWords = 32768;                               // [2][512] u32      =  4 KB
constexpr uint32_t kOffPathS = kOffWords + 4096;                    // [65] uint4
constexpr uint32_t kOffPathT = kOffPathS + 65 * 16;                 // [65] u32
constexpr uint32_t kOffRed = kOffPathT + 65 * 4 + 12;               // [128] u32
constexpr uint32_t kScratchEnd = kOffRed + 512;

enum : int { BAR_TREE = 1, BAR_FULL0 = 2, BAR_EMPTY0 = 4, BAR_SCAN = 6 };

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ uint4 ldg_nc_v4(const uint8_t* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ uint32_t range_mask32(uint64_t rowbase, uint32_t rw, uint64_t first_row,
                                                 uint64_t end_row) {
    if (rowbase >= end_row || rowbase + rw <= first_row) return 0u;
    const uint32_t kmin = rowbase >= first_row ? 0u : (uint32_t)(first_row - rowbase);
    const uint32_t kmax = (rowbase + rw <= end_row) ? rw : (uint32_t)(end_row - rowbase);
    const uint32_t hi = kmax >= 32u ? 0xffffffffu : ((1u << kmax) - 1u);
    const uint32_t lo = kmin >= 32u ? 0xffffffffu : ((1u << kmin) - 1u);
    return hi & ~lo;
}

// ---- scan: selected rows of 32 words -> registers -------------------------------------------------
struct ScanLane {
    uint32_t rpl, h, c;
    const uint8_t* lane_db;  // db + this lane's 16-byte chunk
};

__device__ __forceinline__ ScanLane scan_lane(const AnswerPlan& ap) {
    ScanLane sl;
    const uint32_t lane = threadIdx.x & 31u;
    sl.c = (uint32_t)(ap.row_bytes >> 4);  // 16-byte chunks per row, power of two <= 32
    const uint32_t logc = 31u - __clz(sl.c);
    sl.rpl = 32u >> logc;                  // rows per warp load
    sl.h = lane >> logc;
    sl.lane_db = ap.db + ((lane & (sl.c - 1u)) << 4);
    return sl;
}

// XOR the selected rows of words [g*32, g*32+32) of one task buffer into acc.  `list` is a private
// 2 KB scratch of the calling warp.
template <int UNROLL>
__device__ __forceinline__ void scan_group(const AnswerPlan& ap, const ScanLane& sl, const uint32_t* wbuf,
                                           uint32_t g, uint64_t p, uint16_t* list, uint4& acc) {
    const uint32_t lane = threadIdx.x & 31u, rows_log = ap.D + 1;
    // byte address of task row 0; rows before first_row are never selected
    const int64_t task_row0 = (int64_t)((p << ap.B) << rows_log) - (int64_t)ap.first_row;
    const uint8_t* tbase = sl.lane_db + task_row0 * (int64_t)ap.row_bytes;
    const uint32_t widx = g * 32u + lane;
    uint32_t w = wbuf[widx];
    const uint32_t cnt = __popc(w);
    uint32_t pre = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
        if (lane >= (uint32_t)o) pre += v;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
    uint32_t pos = pre - cnt;
    const uint32_t rbase = widx << rows_log;
    while (w) {
        list[pos++] = (uint16_t)(rbase + (uint32_t)__ffs(w) - 1u);
        w &= w - 1u;
    }
    __syncwarp();
    for (uint32_t e0 = 0; e0 < ((ap.pad & 1u) ? 0u : total); e0 += sl.rpl * UNROLL) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const uint32_t e = e0 + u * sl.rpl + sl.h;
            v[u] = make_uint4(0, 0, 0, 0);
            if (e < total) v[u] = ldg_nc_v4(tbase + (uint64_t)list[e] * ap.row_bytes);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u += 2) {
            acc.x ^= v[u].x ^ v[u + 1].x; acc.y ^= v[u].y ^ v[u + 1].y;
            acc.z ^= v[u].z ^ v[u + 1].z; acc.w ^= v[u].w ^ v[u + 1].w;
        }
    }
    __syncwarp();
}

// fold the sub-rows of a warp load, then XOR into the CTA's shared accumulator
__device__ __forceinline__ void fold_acc(const ScanLane& sl, uint4 acc) {
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t off = 16; off >= sl.c && off > 0; off >>= 1) {
        acc.x ^= __shfl_xor_sync(0xffffffffu, acc.x, off);
        acc.y ^= __shfl_xor_sync(0xffffffffu, acc.y, off);
        acc.z ^= __shfl_xor_sync(0xffffffffu, acc.z, off);
        acc.w ^= __shfl_xor_sync(0xffffffffu, acc.w, off);
    }
    uint32_t* red = reinterpret_cast<uint32_t*>(g_smem2 + kOffRed);
    if (lane < sl.c && (acc.x | acc.y | acc.z | acc.w)) {
        atomicXor(&red[lane * 4 + 0], acc.x);
        atomicXor(&red[lane * 4 + 1], acc.y);
        atomicXor(&red[lane * 4 + 2], acc.z);
        atomicXor(&red[lane * 4 + 3], acc.w);
    }
}

constexpr uint32_t kGroups = kTreeThreads / 32;  // 32-word groups per task

// ---- cooperative node expansion for the narrow top of a task --------------------------------------
// One warp executing a whole 3-block PRG is bound by its own SMSP's ALU rate (~2500 cycles per level).
// Near the root only a few nodes exist, so three warps on three SMSPs each take ONE AES block of the
// same node and write their part of the children straight into shared memory:
//   block 0 -> sL;  block 1 -> tL and bytes 0..14 of sR;  block 2 -> byte 15 of sR and tR.
__device__ __forceinline__ void coop_block(const QueryConst& qc, uint32_t k, uint4 sv, uint32_t t, uint32_t level,
                                           uint32_t lc, uint8_t* left_s, uint8_t* left_t, uint8_t* right_s,
                                           uint8_t* right_t) {
    const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
    uint32_t a[4], n[4], o[4];
#pragma unroll
    for (int c = 0; c < 4; c++) a[c] = s[c] ^ qc.rk[k][c];
#pragma unroll 1
    for (int r = 1; r <= 9; r++) {
        aes_round(a, n, &qc.rk[k][4 * r], lc);
#pragma unroll
        for (int c = 0; c < 4; c++) a[c] = n[c];
    }
#pragma unroll
    for (int c = 0; c < 4; c++) o[c] = aes_last_word(a, c, qc.rk[k][40 + c], s[c], lc);
    const uint32_t m = 0u - t;
    uint32_t cw[4];
#pragma unroll
    for (int c = 0; c < 4; c++) cw[c] = qc.cw[level][c] & m;
    if (k == 0) {
        if (left_s) *reinterpret_cast<uint4*>(left_s) = make_uint4(o[0] ^ cw[0], o[1] ^ cw[1], o[2] ^ cw[2], o[3] ^ cw[3]);
    } else if (k == 1) {
        if (left_t) *left_t = (uint8_t)((o[0] & 1u) ^ (t & (uint32_t)(qc.cw_tl >> level) & 1u));
        if (right_s) {
            uint32_t* rs = reinterpret_cast<uint32_t*>(right_s);
            rs[0] = __funnelshift_r(o[0], o[1], 8) ^ cw[0];
            rs[1] = __funnelshift_r(o[1], o[2], 8) ^ cw[1];
            rs[2] = __funnelshift_r(o[2], o[3], 8) ^ cw[2];
            const uint32_t w3 = (o[3] >> 8) ^ cw[3];
            *reinterpret_cast<uint16_t*>(right_s + 12) = (uint16_t)w3;
            right_s[14] = (uint8_t)(w3 >> 16);
        }
    } else {
        if (right_s) right_s[15] = (uint8_t)(o[0] ^ (cw[3] >> 24));
        if (right_t) *right_t = (uint8_t)(((o[0] >> 8) & 1u) ^ (t & (uint32_t)(qc.cw_tr >> level) & 1u));
    }
}

constexpr uint32_t kCoopMaxLevel = 6;  // breadth-first levels with <= 64 nodes use coop_block

// ---- tree role ------------------------------------------------------------------------------------
__device__ __forceinline__ void tree_role(const QueryConst& qc, const AnswerPlan& ap, uint32_t lc) {
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t D = ap.D, B = ap.B, Fc = ap.Fc;
    const uint32_t lvl0 = Fc + B;
    const uint32_t RW = 2u << D;
    const uint32_t nlp = 1u << D;
    const uint32_t rows_log = D + 1;
    const uint64_t end_row = ap.first_row + ap.n_rows;
    uint32_t* stk = reinterpret_cast<uint32_t*>(g_smem2 + kOffStack);
    uint4* bfs_s = reinterpret_cast<uint4*>(g_smem2 + kOffBfsS);
    uint8_t* bfs_t = g_smem2 + kOffBfsT;
    uint32_t* words = reinterpret_cast<uint32_t*>(g_smem2 + kOffWords);
    uint4* path_s = reinterpret_cast<uint4*>(g_smem2 + kOffPathS);
    uint8_t* path_t = g_smem2 + kOffPathT;

    // contiguous block of tasks for this CTA
    const uint64_t k0 = (uint64_t)blockIdx.x * ap.task_count / gridDim.x;
    const uint64_t k1 = (uint64_t)(blockIdx.x + 1) * ap.task_count / gridDim.x;
    uint64_t p_prev = 0;
    for (uint64_t k = k0; k < k1; k++) {
        const uint32_t it = (uint32_t)(k - k0), buf = it & 1u;
        const uint64_t p = ap.task_first + k;  // node index at level Fc
        // -- top walk: re-expand from the common ancestor with the previous task's root (3 threads) --
        const uint32_t ck = warp & 3u, cj = (warp >> 2) * 32u + (tid & 31u);  // coop role: block ck of node cj
        const bool coop_warp = ck < 3u && warp < 8u;
        {
            uint32_t cp = 0;
            if (it == 0) {
                if (tid == 0) {
                    path_s[0] = make_uint4(qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]);
                    path_t[0] = (uint8_t)qc.t_init;
                }
                bar_sync(BAR_TREE, kTreeThreads);
            } else {
                cp = Fc - (64u - (uint32_t)__clzll((long long)(p ^ p_prev)));
            }
            for (uint32_t l = cp; l < Fc; l++) {
                if (coop_warp && cj == 0) {
                    const uint32_t bit = (uint32_t)(p >> (Fc - 1 - l)) & 1u;
                    uint8_t* ds = reinterpret_cast<uint8_t*>(&path_s[l + 1]);
                    coop_block(qc, ck, path_s[l], path_t[l], l, lc, bit ? nullptr : ds, bit ? nullptr : &path_t[l + 1],
                               bit ? ds : nullptr, bit ? &path_t[l + 1] : nullptr);
                }
                bar_sync(BAR_TREE, kTreeThreads);
            }
            if (tid == 0) {
                bfs_s[0] = path_s[Fc];
                bfs_t[0] = path_t[Fc];
            }
        }
        p_prev = p;
        bar_sync(BAR_TREE, kTreeThreads);
        // -- B breadth-first levels: 2^B thread roots --
        for (uint32_t b = 0; b < B; b++) {
            const uint32_t o = ((b + 1) & 1) * kTreeThreads;
            if (b <= kCoopMaxLevel) {
                if (coop_warp && cj < (1u << b)) {
                    uint8_t* cs = reinterpret_cast<uint8_t*>(&bfs_s[o + 2 * cj]);
                    coop_block(qc, ck, bfs_s[(b & 1) * kTreeThreads + cj], bfs_t[(b & 1) * kTreeThreads + cj], Fc + b, lc,
                               cs, &bfs_t[o + 2 * cj], cs + 16, &bfs_t[o + 2 * cj + 1]);
                }
            } else if (tid < (1u << b)) {
                const uint4 sv = bfs_s[(b & 1) * kTreeThreads + tid];
                const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
                Children ch;
                expand_node<1>(qc, s, bfs_t[(b & 1) * kTreeThreads + tid], Fc + b, lc, ch);
                bfs_s[o + 2 * tid] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                bfs_s[o + 2 * tid + 1] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                bfs_t[o + 2 * tid] = (uint8_t)ch.tl;
                bfs_t[o + 2 * tid + 1] = (uint8_t)ch.tr;
            }
            bar_sync(BAR_TREE, kTreeThreads);
        }
        uint32_t s[4] = {0, 0, 0, 0}, t = 0;
        const bool active = tid < (1u << B);
        if (active) {
            const uint4 sv = bfs_s[(B & 1) * kTreeThreads + tid];
            s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
            t = bfs_t[(B & 1) * kTreeThreads + tid];
        }
        bar_sync(BAR_TREE, kTreeThreads);  // the sibling stack aliases the BFS buffers

        const uint64_t f = (p << B) + tid;  // this thread's subtree root at level lvl0
        const uint64_t rowbase = f << rows_log;
        const uint64_t warp_row0 = ((p << B) + (tid & ~31u)) << rows_log;
        const uint64_t warp_row1 = ((p << B) + (tid & ~31u) + 32u) << rows_log;
        uint32_t word = 0;
        const bool warp_live = (warp * 32u < (1u << B)) && warp_row0 < end_row && warp_row1 > ap.first_row;
        if (warp_live) {
            uint32_t tst = 0, d = 0, i = 0, bw = 0;
            while (true) {
                const uint32_t lvl = lvl0 + d;
                const uint32_t cw[4] = {qc.cw[lvl][0], qc.cw[lvl][1], qc.cw[lvl][2], qc.cw[lvl][3]};
                const uint32_t cwtl = (uint32_t)(qc.cw_tl >> lvl) & 1u, cwtr = (uint32_t)(qc.cw_tr >> lvl) & 1u;
                uint32_t a[3][4];
                aes3_main_rounds(qc, s, lc, a);
                if (d < D) {
                    Children ch;
                    inner_tail(qc, a, s, t, cw, cwtl, cwtr, lc, ch);
#pragma unroll
                    for (int w = 0; w < 4; w++) stk[(d * 4 + w) * kTreeThreads + tid] = ch.sr[w];
                    tst = (tst & ~(1u << d)) | (ch.tr << d);
#pragma unroll
                    for (int w = 0; w < 4; w++) s[w] = ch.sl[w];
                    t = ch.tl;
                    d++;
                    continue;
                }
                bw |= leaf_parent_tail(qc, a, s, t, cw, cwtl, cwtr, lc) << (2u * i);
                i++;
                if (i == nlp) break;
                d = D - (uint32_t)(__ffs(i) - 1);  // next subtree root: the sibling stacked at that depth
#pragma unroll
                for (int w = 0; w < 4; w++) s[w] = stk[((d - 1) * 4 + w) * kTreeThreads + tid];
                t = (tst >> (d - 1)) & 1u;
            }
            word = active ? (bw & range_mask32(rowbase, RW, ap.first_row, end_row)) : 0u;
        }
        if (it >= 2) bar_sync(BAR_EMPTY0 + buf, kAnswerThreads);  // scan warps are done with this buffer
        words[buf * kTreeThreads + tid] = word;
        __threadfence_block();
        bar_arrive(BAR_FULL0 + buf, kAnswerThreads);
        if (k + 1 == k1) {
            // last task of this CTA: nothing left to expand, so the tree warps scan groups
            // [kScanWarps, kGroups) themselves (row lists in the now idle sibling stack)
            bar_sync(BAR_TREE, kTreeThreads);
            const ScanLane sl = scan_lane(ap);
            uint4 acc = make_uint4(0, 0, 0, 0);
            uint16_t* list = reinterpret_cast<uint16_t*>(g_smem2 + kOffStack) + warp * kListEntries;
            for (uint32_t g = kScanWarps + warp; g < kGroups; g += kTreeThreads / 32)
                scan_group<8>(ap, sl, words + buf * kTreeThreads, g, p, list, acc);
            fold_acc(sl, acc);
        }
    }
    // drain the scan warps' last "empty" arrivals so that no barrier is left half-complete
    const uint32_t ntask = (uint32_t)(k1 - k0);
    if (ntask >= 2) bar_sync(BAR_EMPTY0 + (ntask & 1u), kAnswerThreads);
    if (ntask >= 1) bar_sync(BAR_EMPTY0 + ((ntask - 1) & 1u), kAnswerThreads);
}

// ---- scan role ------------------------------------------------------------------------------------
__device__ __forceinline__ void scan_role(const AnswerPlan& ap) {
    const uint32_t sw = (threadIdx.x - kTreeThreads) >> 5;
    const ScanLane sl = scan_lane(ap);
    const uint32_t* words = reinterpret_cast<const uint32_t*>(g_smem2 + kOffWords);
    const uint32_t dyn_base = smem_addr(g_smem2);
    uint16_t* list = reinterpret_cast<uint16_t*>(g_smem2 + (kTableEnd - dyn_base)) + sw * kListEntries;  // above the tables
    uint4 acc = make_uint4(0, 0, 0, 0);
    const uint64_t k0 = (uint64_t)blockIdx.x * ap.task_count / gridDim.x;
    const uint64_t k1 = (uint64_t)(blockIdx.x + 1) * ap.task_count / gridDim.x;
    for (uint64_t k = k0; k < k1; k++) {
        const uint32_t buf = (uint32_t)(k - k0) & 1u;
        bar_sync(BAR_FULL0 + buf, kAnswerThreads);
        // the tree warps help with the CTA's last task (they have nothing left to expand): groups
        // [0, kScanWarps) stay here, the rest is theirs
        const uint32_t gend = (k + 1 == k1) ? (uint32_t)kScanWarps : kGroups;
        for (uint32_t g = sw; g < gend; g += kScanWarps)
            scan_group<16>(ap, sl, words + buf * kTreeThreads, g, ap.task_first + k, list, acc);
        bar_arrive(BAR_EMPTY0 + buf, kAnswerThreads);
    }
    fold_acc(sl, acc);
}

__global__ void __launch_bounds__(kAnswerThreads, 1)
answer_kernel(const __grid_constant__ QueryConst qc, const __grid_constant__ AnswerPlan ap,
              const uint32_t* __restrict__ t0) {
    {
        const uint32_t base = smem_addr(g_smem2);
        uint32_t dyn;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
        if (base + dyn < kTableEnd + kListBytes || base + kScratchEnd > kTableBase) __trap();
    }
    fill_tables(t0);
    uint32_t* red = reinterpret_cast<uint32_t*>(g_smem2 + kOffRed);
    if (threadIdx.x < 128) red[threadIdx.x] = 0;
    __syncthreads();
    if (threadIdx.x < kTreeThreads)
        tree_role(qc, ap, lane_const());
    else
        scan_role(ap);
    __syncthreads();
    const uint32_t words_out = (uint32_t)(ap.row_bytes >> 2);
    if (threadIdx.x < words_out) {
        const uint32_t v = red[threadIdx.x];
        if (v) atomicXor(&ap.out[threadIdx.x], v);
    }
}

bool answer_kernel_supports(uint64_t row_bytes, uint32_t D) {
    if (row_bytes < 16 || row_bytes > 512 || (row_bytes & (row_bytes - 1))) return false;
    return D <= (uint32_t)kV2MaxDepth;
}

cudaError_t configure_answer_kernel() {
    return cudaFuncSetAttribute(answer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(kTableEnd + kListBytes));
}

cudaError_t launch_answer(const QueryConst& qc, const AnswerPlan& ap, const uint32_t* d_t0, uint32_t smem_base,
                          int num_sms, cudaStream_t st) {
    if (ap.task_count == 0) return cudaSuccess;
    unsigned grid = (unsigned)(ap.task_count < (uint64_t)num_sms ? ap.task_count : (uint64_t)num_sms);
    answer_kernel<<<grid, kAnswerThreads, kTableEnd + kListBytes - smem_base, st>>>(qc, ap, d_t0);
    count_launch();
    return cudaGetLastError();
}

}  // namespace pirb
