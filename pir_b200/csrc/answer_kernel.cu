// answer_kernel.cu — the dominant kernel: full-domain DPF evaluation fused with the masked row XOR
// (db.go:68-174 + dpf/server.go:55-101 + slot.go:19-30), warp-specialised for sm_100a.
//
// One persistent CTA per SM, 16 "tree" warps + kScanWarps "scan" warps:
//   tree warps   per CTA task (512 thread roots over consecutive rows, TaskShape below): expand the node
//                range above the roots breadth-first from the tree root, then every thread walks its own
//                2^(d+1)-row subtree depth-first (sibling stack in shared memory, d <= 7) with the
//                unrolled T-table AES of aes_unrolled.cuh.  Selection bits are published in SLICES of 8
//                rows per thread into a ring of word buffers in shared memory.  Tree warps never touch
//                global memory in the steady state.
//   scan warps   turn a slice's words into a compact list of selected rows (popc prefix sum per 32
//                words) and stream exactly those rows from HBM with coalesced 128-bit loads (32/c rows
//                per warp load for rows of c 16-byte chunks), XOR-accumulating in registers; rows whose
//                bit is 0 are never fetched (like db.go:93).  Loads go straight to registers: the
//                shared-memory pipe is the kernel's bottleneck (AES table lookups), so staging rows
//                through shared memory (TMA / cp.async) would cost more than it hides.
// Hand-off is a producer/consumer ring of kRing slice buffers guarded by mbarriers (full: 16 tree warps
// arrive, empty: the scan warps arrive).  The answer is reduced per CTA and XORed into global memory
// (possibly a peer GPU's) with one atomic per word.  When a CTA has several tasks, the last scan warp also
// walks the narrow top of the tasks ahead (TopWalker) so that the tree warps find it ready.
//
// Why deep tasks: the top of a task is executed by few warps and is bound by single-warp issue latency
// (~1.7k cycles per level, measured), during which the lookup pipe idles.  A thread that owns 2^(d+1)
// rows amortises it over 2^(d+1)-1 depth-first nodes.
#include "kernels.cuh"
#include "aes_unrolled.cuh"

namespace pirb {

extern __shared__ __align__(16) uint8_t g_smem2[];

constexpr int kTreeThreads = 512;
constexpr int kTreeWarps = kTreeThreads / 32;
#ifndef PIRB_SCAN_WARPS
#define PIRB_SCAN_WARPS 4
#endif
constexpr int kScanWarps = PIRB_SCAN_WARPS;
constexpr int kConsumerWarps = kTreeWarps + kScanWarps;        // warps that may run consume_slices
constexpr int kAnswerThreads = 32 * kConsumerWarps;            // (the last scan warp also walks the tops of the tasks ahead)
constexpr uint32_t kHand = 128;                // nodes handed from the top warp to the tree warps per task
constexpr int kV2MaxDepth = 7;                 // a thread owns at most 2^8 rows (sibling stack: 7 x 16 B)
constexpr uint32_t kSliceLog = 3;              // rows per thread per slice: at most 8; AnswerPlan::slog_cap picks 4 for shallow tasks
constexpr uint32_t kRing = 8;                  // slice buffers in flight between tree and scan warps
constexpr uint32_t kListEntries = 256;         // rows selected out of 32 words of a slice, per warp

// scratch below the tables (offsets from the dynamic shared memory base)
constexpr uint32_t kOffStack = 0;                                   // [D<=7][4][512] u32 = 56 KB (aliases BFS)
constexpr uint32_t kOffBfsS = 0;                                    // [2][512] uint4    = 16 KB
constexpr uint32_t kOffBfsT = 16384;                                // [2][512] u8       =  1 KB
constexpr uint32_t kMaxTasks = 512;            // per CTA (launch_answer sizes the grid accordingly)
constexpr uint32_t kOffPathS = kV2MaxDepth * 4 * kTreeThreads * 4;  // [65] uint4: the single-node levels above a task
constexpr uint32_t kOffPathT = kOffPathS + 65 * 16;                 // [65] u8 (+ pad)
constexpr uint32_t kOffTasks = kOffPathT + 80;                      // task list: header (n, base, stride) + [kMaxTasks] u32
constexpr uint32_t kTaskHeader = 8;                                 // u32 words before the entries
constexpr uint32_t kOffTopScrS = (kOffTasks + 4 * (kTaskHeader + kMaxTasks) + 15u) & ~15u;  // top warp: [kHand] uint4 level buffer
constexpr uint32_t kOffTopScrT = kOffTopScrS + kHand * 16;          // [kHand] u8
constexpr uint32_t kOffPath2S = kOffTopScrT + kHand;                // top warp: [65] uint4 single-node levels
constexpr uint32_t kOffPath2T = kOffPath2S + 65 * 16;               // [65] u8 (+ pad)
constexpr uint32_t kOffTopState = kOffPath2T + 80;                  // top walker: next task, Ls of the previous one, its first row
constexpr uint32_t kOffRk = kOffTopState + 16;                      // [3][44] u32: the round keys, for quad_path_walk
constexpr uint32_t kScratchEnd = kOffRk + 3 * 44 * 4;
// above the tables (absolute shared-window addresses)
constexpr uint32_t kAbsLists = kTableEnd;                                      // [20 warps][256] u16
constexpr uint32_t kAbsRing = kAbsLists + kConsumerWarps * kListEntries * 2;    // [kRing][512] u32
constexpr uint32_t kAbsBars = kAbsRing + kRing * kTreeThreads * 4;             // full[kRing], empty[kRing] u64
constexpr uint32_t kAbsClaim = kAbsBars + 2 * kRing * 8;                       // claim counter, done[kRing] u32
constexpr uint32_t kAbsRed = kAbsClaim + 4 * (1 + kRing) + 12;                 // [kMaxRowBytes / 4] u32: the CTA's partial answer
constexpr uint32_t kMaxRowBytes = 4096;                                        // rows of 16..4096 bytes (power of two)
constexpr uint32_t kAbsHandS = kAbsRed + kMaxRowBytes;                          // [2][kHand] uint4: top warp -> tree warps
constexpr uint32_t kAbsHandT = kAbsHandS + 2 * kHand * 16;                     // [2][kHand] u8
constexpr uint32_t kAbsTopBars = kAbsHandT + 2 * kHand;                        // hand_full[2], hand_empty[2] u64
constexpr uint32_t kSmemEnd = kAbsTopBars + 4 * 8;
static_assert(kScratchEnd <= kTableBase, "scratch below the tables overflows into them");
static_assert(kSmemEnd <= 227u * 1024u, "more shared memory than an sm_100 CTA can have");
__device__ __forceinline__ uint32_t* red_words() {
    return reinterpret_cast<uint32_t*>(g_smem2 + (kAbsRed - smem_addr(g_smem2)));
}

enum : int { BAR_TREE = 1 };

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ---- mbarrier ring ----------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t addr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t addr) {
    asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(addr), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t full_bar(uint32_t slot) { return kAbsBars + slot * 8u; }
__device__ __forceinline__ uint32_t empty_bar(uint32_t slot) { return kAbsBars + (kRing + slot) * 8u; }
__device__ __forceinline__ uint32_t* ring_words(uint32_t slot) {
    return reinterpret_cast<uint32_t*>(g_smem2 + (kAbsRing - smem_addr(g_smem2))) + slot * kTreeThreads;
}

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

// rows per thread per slice (log2) for a task of depth d: 8, or all 2^(d+1) rows when there are fewer
__device__ __forceinline__ uint32_t slice_log(uint32_t d, uint32_t cap) { return d + 1u < cap ? d + 1u : cap; }

// ---- scan: selected rows of 32 words of a slice -> registers ----------------------------------------
// Rows of any multiple of 16 bytes up to 4 KB.  A row of c 16-byte chunks takes c lanes of a warp load; rpl = the largest
// power of two <= 32 / c rows are fetched per load (the list is kept as rpl interleaved sub-lists) and the lanes beyond
// rpl * c idle: none for power-of-two rows, 11 of 32 for 336-byte rows (c = 21), 8 of 32 for 48-byte rows (c = 3,
// rpl = 8).  Rows wider than a warp load (512 B) are fetched in NCB column blocks of 512 B, the last one partly filled.
struct ScanLane {
    uint32_t rpl, h, c, chunk;
    uint32_t hsub;           // the sub-list this lane reads its row numbers from
    bool live;               // this lane fetches (lanes beyond rpl * c idle)
    const uint8_t* lane_db;  // db + this lane's 16-byte chunk
};

__device__ __forceinline__ ScanLane scan_lane(const AnswerPlan& ap) {
    ScanLane sl;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t c_row = (uint32_t)(ap.row_bytes >> 4);  // 16-byte chunks per row
    sl.c = c_row > 32u ? 32u : c_row;
    sl.rpl = 1u << (31u - __clz(32u / sl.c));              // rows per warp load
    sl.h = lane / sl.c;
    sl.chunk = lane - sl.h * sl.c;
    sl.live = sl.h < sl.rpl;
    sl.hsub = sl.live ? sl.h : sl.rpl - 1u;                // (idle lanes read a valid sub-list ...
    if (!sl.live) sl.h = 0x10000u;                         //  ... and their entry numbers lie beyond every list: no load)
    sl.lane_db = ap.db + (sl.chunk << 4);
    return sl;
}

// XOR the selected rows of words [g*32, g*32+32) of one slice buffer into acc.  Word `i` of the buffer
// holds thread i's selection bits for its rows [slice_off, slice_off + 2^slice_log) (lane l of the group
// owns rows [group_row0 + (l << rows_log), + 2^rows_log), numbered from the shard's first row).  `list` is
// a private 512-byte scratch of the calling warp.
template <int UNROLL, int NCB>
__device__ __forceinline__ void scan_group(const AnswerPlan& ap, const ScanLane& sl, const uint32_t* wbuf,
                                           uint32_t g, int64_t group_row0, uint32_t rows_log, uint32_t slice_off,
                                           uint16_t* list, uint4 (&acc)[NCB]) {
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t w = wbuf[g * 32u + lane];
    const uint32_t cnt = __popc(w);
    uint32_t pre = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
        if (lane >= (uint32_t)o) pre += v;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
    if (total == 0) return;
    // byte address of the group's first row, for this lane's 16-byte chunk (rows before the shard's
    // first row are never selected, so a negative offset is never dereferenced)
    const uint8_t* gbase = sl.lane_db + group_row0 * (int64_t)ap.row_bytes;
#ifdef PIRB_TEX_ROWS
    // texel (16-byte chunk) index of this lane's chunk of the group's first row; rows of the shard lie below 2^28 texels
    const int tbase = (int)((int64_t)sl.chunk + group_row0 * (int64_t)(ap.row_bytes >> 4));
    const int trow = (int)(ap.row_bytes >> 4);
#endif
    // The list is kept as rpl interleaved sub-lists (entry p goes to sub-list p % rpl, the lane group that
    // will fetch it), so a lane group reads its next 8 row numbers with ONE 128-bit shared load.
    uint32_t pos = pre - cnt;
    const uint32_t rbase = (lane << rows_log) + slice_off;
    const uint32_t lrpl = 31u - __clz(sl.rpl), sub_stride = kListEntries >> lrpl;
    while (w) {
        list[(pos & (sl.rpl - 1u)) * sub_stride + (pos >> lrpl)] = (uint16_t)(rbase + (uint32_t)__ffs(w) - 1u);
        pos++;
        w &= w - 1u;
    }
    __syncwarp();
    const uint16_t* sub = list + sl.hsub * sub_stride;
#pragma unroll
    for (int cb = 0; cb < NCB; cb++) {
        const uint8_t* cbase = gbase + cb * 512;
        // wide rows: the last 512-byte column block may be partly filled (lanes beyond the row's last chunk skip it)
        uint32_t my_total = total;
        if (NCB > 1) {
            const uint32_t c_row = (uint32_t)(ap.row_bytes >> 4);
            if ((uint32_t)cb * 32u >= c_row) break;
            if ((uint32_t)cb * 32u + sl.chunk >= c_row) my_total = 0u;
        }
        for (uint32_t e0 = 0; e0 < ((PIRB_AP_PAD(ap) & 1u) ? 0u : total); e0 += sl.rpl * UNROLL) {
            uint4 v[UNROLL];
#pragma unroll
            for (int k = 0; k < UNROLL / 8; k++) {
                const uint32_t eb = e0 + (uint32_t)k * 8u * sl.rpl;  // first entry of this block of 8 loads
                uint4 L = make_uint4(0, 0, 0, 0);
                if (eb < total) L = *reinterpret_cast<const uint4*>(sub + (eb >> lrpl));
                const uint32_t lw[4] = {L.x, L.y, L.z, L.w};
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const uint32_t e = eb + (uint32_t)u * sl.rpl + sl.h;
                    const uint32_t row = (lw[u >> 1] >> ((u & 1) * 16)) & 0xffffu;
                    v[k * 8 + u] = make_uint4(0, 0, 0, 0);
#ifdef PIRB_TEX_ROWS
                    if (e < my_total) v[k * 8 + u] = tex1Dfetch<uint4>((cudaTextureObject_t)ap.tex, tbase + cb * 32 + (int)row * trow);
#else
                    if (e < my_total) v[k * 8 + u] = ldg_nc_v4(cbase + (uint64_t)row * ap.row_bytes);
#endif
                }
            }
#pragma unroll
            for (int u = 0; u < UNROLL; u += 2) {
                acc[cb].x ^= v[u].x ^ v[u + 1].x; acc[cb].y ^= v[u].y ^ v[u + 1].y;
                acc[cb].z ^= v[u].z ^ v[u + 1].z; acc[cb].w ^= v[u].w ^ v[u + 1].w;
            }
        }
    }
    __syncwarp();
}

// fold the sub-rows of a warp load, then XOR into the CTA's shared accumulator
template <int NCB>
__device__ __forceinline__ void fold_acc(const ScanLane& sl, uint4 (&accs)[NCB]) {
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t* red = red_words();
    const bool pow2 = (sl.c & (sl.c - 1u)) == 0u;
#pragma unroll
    for (int cb = 0; cb < NCB; cb++) {
        uint4 acc = accs[cb];
        if (pow2) {  // lanes l, l + c, l + 2c ... hold the same chunk: butterfly, then one atomic per chunk
            for (uint32_t off = 16; off >= sl.c && off > 0; off >>= 1) {
                acc.x ^= __shfl_xor_sync(0xffffffffu, acc.x, off);
                acc.y ^= __shfl_xor_sync(0xffffffffu, acc.y, off);
                acc.z ^= __shfl_xor_sync(0xffffffffu, acc.z, off);
                acc.w ^= __shfl_xor_sync(0xffffffffu, acc.w, off);
            }
        }
        // (other row sizes: every fetching lane adds its own share; the accumulators of idle lanes are zero)
        if ((pow2 ? lane < sl.c : sl.live) && (acc.x | acc.y | acc.z | acc.w)) {
            atomicXor(&red[cb * 128 + sl.chunk * 4 + 0], acc.x);
            atomicXor(&red[cb * 128 + sl.chunk * 4 + 1], acc.y);
            atomicXor(&red[cb * 128 + sl.chunk * 4 + 2], acc.z);
            atomicXor(&red[cb * 128 + sl.chunk * 4 + 3], acc.w);
        }
    }
}

constexpr uint32_t kGroups = kTreeThreads / 32;  // 32-word groups per slice

// level-L nodes (2^(D+1) rows each) a uniform task spans: the shape's own 32*hiw + 16*(16 - hiw), or fewer when the
// host trims the tasks so that they fill every SM equally (AnswerPlan::m_task; the trailing threads then idle)
__host__ __device__ __forceinline__ uint32_t task_nodes(uint32_t hi_warps, uint32_t m_task) {
    const uint32_t nom = 32u * hi_warps + 16u * (16u - hi_warps);
    return (m_task != 0u && m_task < nom) ? m_task : nom;
}
__host__ __device__ __forceinline__ uint64_t answer_tasks_of(uint64_t first_row, uint64_t n_rows, uint32_t D, uint32_t hi_warps,
                                                             uint32_t m_task) {
    const uint64_t R = (uint64_t)task_nodes(hi_warps, m_task) << (D + 1u);
    const uint64_t T0 = first_row & ~((2ull << D) - 1ull);
    return (first_row + n_rows - T0 + R - 1ull) / R;
}

// ---- work split ---------------------------------------------------------------------------------------
// A TASK is 512 thread roots over consecutive rows, of shape (D, hi_warps): the threads of the first
// hi_warps tree warps own 2^(D+1) rows each, those of the other 16 - hi_warps warps half as many (their roots
// sit one tree level lower), R = 2^(D+5) * (16 + hi_warps) rows in all: 512 * 2^(D+1) times 1, 7/8, 3/4 or
// 5/8.  hi_warps is a multiple of 4: warp w runs on SM sub-partition w % 4 and the table-lookup rate is
// capped per sub-partition (a CTA with 13..16 equally loaded warps takes the same time, measured), so
// every sub-partition gets the same mix.  Tasks need not be aligned subtrees: the top of a task is a
// breadth-first expansion of the node RANGE above its thread roots (tree_role).  Two ways to cover a shard,
// chosen by the host against a cost model (capi.cu choose_answer_split):
//   uniform (ap.hi_warps != 0)  every task has the shape (ap.D, ap.hi_warps); task t covers rows
//            [T0 + t*R, T0 + (t+1)*R), T0 = first_row rounded down to a thread root, and CTA b runs tasks
//            b, b + gridDim.x, ...  With R close to shard_rows / num_sms this is ONE task (one task top)
//            per CTA on every SM: the split of choice for shards of a multi-GPU server.
//   binary (ap.hi_warps == 0)   the shard is cut into CHUNKS of 2^(10+Dmin) rows, CTA b owns the contiguous
//            chunk range [b*C/G, (b+1)*C/G) and covers it with aligned power-of-two tasks (shape (d, 16),
//            Dmin <= d <= ap.D, as large as alignment allows), smallest first.  All 16 warps of a task do
//            the same work, which the latency-sensitive even-FinalCW path likes: the split of choice when
//            the uniform one would leave warps idle half of the time (large single-GPU databases).
// Thread kAnswerThreads-1 writes the CTA's task list to shared memory once; both roles walk it.
struct TaskShape {
    uint32_t D, hiw;       // depth of the hi threads' subtrees; number of hi warps (4, 8, 12 or 16)
    uint32_t m_hi, m_nom;  // hi thread roots; nodes of the hi roots' tree level the task spans (hi roots + lo parents)
    uint32_t slog, n_slices_log;
    uint64_t R;
    __device__ __forceinline__ void set(uint32_t d, uint32_t hi_warps, uint32_t slog_cap, uint32_t m_task) {
        D = d;
        hiw = hi_warps;
        m_hi = 32u * hiw;
        m_nom = task_nodes(hiw, m_task);
        slog = slice_log(D, slog_cap);
        n_slices_log = D + 1u - slog;
        R = (uint64_t)m_nom << (D + 1u);
    }
    // first row, log2(rows per thread) and log2(rows per thread per slice) of tree warp g's threads
    __device__ __forceinline__ void group_geom(uint32_t g, uint64_t row0, uint64_t& grow0, uint32_t& rows_log, uint32_t& gslog) const {
        if (g < hiw) {
            grow0 = row0 + ((uint64_t)(g * 32u) << (D + 1u));
            rows_log = D + 1u;
            gslog = slog;
        } else {
            grow0 = row0 + ((uint64_t)m_hi << (D + 1u)) + ((uint64_t)((g - hiw) * 32u) << D);
            rows_log = D;
            gslog = slog - 1u;
        }
    }
};

// number of tree levels above nodes [a, a + m) of level L that hold a single node (the common prefix of a and a+m-1)
__device__ __forceinline__ uint32_t single_node_levels(uint32_t L, uint64_t a, uint32_t m) {
    return L - (64u - (uint32_t)__clzll((long long)(a ^ (a + m - 1u))));  // (clzll(0) = 64)
}

// geometry of the binary split
struct ChunkRange {
    uint32_t chunk_log;
    uint64_t c0, c1;  // this CTA's chunks
    __device__ __forceinline__ ChunkRange(const AnswerPlan& ap) {
        chunk_log = 10u + ap.Dmin;
        const uint64_t end_row = ap.first_row + ap.n_rows;
        const uint64_t cf = ap.first_row >> chunk_log, cl = (end_row + ((1ull << chunk_log) - 1ull)) >> chunk_log;
        c0 = cf + (uint64_t)blockIdx.x * (cl - cf) / gridDim.x;
        c1 = cf + (uint64_t)(blockIdx.x + 1) * (cl - cf) / gridDim.x;
    }
};

// The CTA's task list in shared memory: n, the slice cap (log2), then (64-bit) base and stride, m_task, then n entries
// (offset << 5 | (hi_warps/4 - 1) << 3 | d); a task's first row is base + offset * stride.
struct TaskIter {
    const uint32_t* list;
    uint32_t n, pos;
    __device__ __forceinline__ TaskIter(const uint32_t* task_list) : list(task_list), n(task_list[0]), pos(0) {}
    __device__ __forceinline__ void decode(uint32_t k, uint64_t& row_start, TaskShape& ts) const {
        const uint32_t e = list[kTaskHeader + k];
        const uint64_t base = *reinterpret_cast<const uint64_t*>(list + 2), stride = *reinterpret_cast<const uint64_t*>(list + 4);
        ts.set(e & 7u, 4u * (((e >> 3) & 3u) + 1u), list[1], list[6]);
        row_start = base + (uint64_t)(e >> 5) * stride;
    }
    __device__ __forceinline__ uint32_t slices_log(uint32_t k) const {  // log2(slices of task k)
        const uint32_t d = list[kTaskHeader + k] & 7u;
        return d + 1u - slice_log(d, list[1]);
    }
};

__device__ __forceinline__ void build_task_list(const AnswerPlan& ap, uint32_t* task_list) {
    uint32_t n = 0;
    uint64_t base, stride;
    uint32_t* entries = task_list + kTaskHeader;
    if (ap.hi_warps) {
        TaskShape ts;
        ts.set(ap.D, ap.hi_warps, ap.slog_cap, ap.m_task);
        base = (ap.first_row & ~((2ull << ap.D) - 1ull)) + (uint64_t)blockIdx.x * ts.R;
        stride = (uint64_t)gridDim.x * ts.R;
        const uint64_t NT = answer_tasks_of(ap.first_row, ap.n_rows, ap.D, ap.hi_warps, ap.m_task);
        if (blockIdx.x < NT) n = (uint32_t)((NT - blockIdx.x + gridDim.x - 1ull) / gridDim.x);
        if (n > kMaxTasks) __trap();  // launch_answer sizes the grid so that this cannot happen
        for (uint32_t k = 0; k < n; k++) entries[k] = (k << 5) | ((ap.hi_warps / 4u - 1u) << 3) | ap.D;
    } else {
        const ChunkRange cr(ap);
        base = cr.c0 << cr.chunk_log;
        stride = 1ull << cr.chunk_log;
        for (uint32_t want = ap.Dmin; want <= ap.D; want++) {  // smallest tasks first: whatever runs last is the
            uint64_t cur = cr.c0;                             // part whose scan is overlapped least
            while (cur < cr.c1) {
                uint32_t d = ap.Dmin;
                uint64_t span = 1;
                while (d < ap.D && !(cur & (2ull * span - 1ull)) && cur + 2ull * span <= cr.c1) {
                    d++;
                    span *= 2ull;
                }
                if (d == want) {
                    if (n >= kMaxTasks) __trap();
                    entries[n++] = ((uint32_t)(cur - cr.c0) << 5) | (3u << 3) | d;
                }
                cur += span;
            }
        }
    }
    task_list[0] = n;
    task_list[1] = ap.slog_cap;
    task_list[6] = ap.hi_warps ? ap.m_task : 0u;
    *reinterpret_cast<uint64_t*>(task_list + 2) = base;
    *reinterpret_cast<uint64_t*>(task_list + 4) = stride;
}

// ---- cooperative node expansion for the narrow top of a task --------------------------------------
// One warp executing a whole 3-block PRG is bound by its own SMSP's ALU rate (~2500 cycles per level).
// Near the root only a few nodes exist, so three warps on three SMSPs each take ONE AES block of the
// same node and write their part of the children straight into shared memory:
//   block 0 -> sL;  block 1 -> tL and bytes 0..14 of sR;  block 2 -> byte 15 of sR and tR.
template <class QC>
__device__ __forceinline__ void coop_block(const QC& qc, uint32_t k, uint4 sv, uint32_t t, uint32_t level,
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


// ---- the single-node levels, by twelve lanes of one warp -------------------------------------------------
// The levels above a task that hold a single node are one chain of dependent AES calls: what counts is latency.
// Lane 4*k + c (k = 0..2, c = 0..3) computes column c of PRG block k: 3 SHFL (the other columns of the previous
// round) + 4 PRMT + 4 LDS + 2 LOP3 per round, 86 cycles against 135 for a thread that computes a whole block
// (profiles/r1_latency_probe.txt).  The node stays in registers from level to level (the chosen child's words
// come back by SHFL), because handing it over through shared memory costs ~600 cycles per level, as much as six
// rounds; every level is also stored to path_s/path_t, off the critical path, for the restart of later tasks.
// Round keys come from shared memory (conflict-free: the twelve words of a round lie in twelve banks).
// Levels [l0, l1) of the walk towards node `a` of level L; path_s[l0], path_t[l0] hold the start node.
// (QC: QueryConst, or the LaConst of the NEXT query for the look-ahead; rk_base: the [3][44] round keys in shared memory)
template <class QC>
__device__ __forceinline__ void quad_path_walk(const QC& qc, const uint32_t* rk_base, uint32_t lc, uint64_t a, uint32_t L, uint32_t l0,
                                               uint32_t l1, uint4* path_s, uint8_t* path_t, unsigned long long* dbg = nullptr) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t c = lane & 3u, k = (lane >> 2) % 3u, g0 = lane & ~3u;  // (lanes 12..31 run along and store nothing)
    const uint32_t* rk = rk_base + k * 44u + c;
    uint32_t sw = reinterpret_cast<const uint32_t*>(&path_s[l0])[c];
    uint32_t t = path_t[l0];
    for (uint32_t l = l0; l < l1; l++) {
        const uint32_t cwc = qc.cw[l][c] & (0u - t);
        const uint32_t bit = (uint32_t)(a >> (L - 1u - l)) & 1u;
        uint32_t x = sw ^ rk[0];
        uint32_t x1, x2, x3;
#pragma unroll 1
        for (int r = 1; r <= 9; r++) {
            x1 = __shfl_sync(0xffffffffu, x, g0 + ((c + 1u) & 3u));
            x2 = __shfl_sync(0xffffffffu, x, g0 + ((c + 2u) & 3u));
            x3 = __shfl_sync(0xffffffffu, x, g0 + ((c + 3u) & 3u));
            const uint32_t v0 = lds0(__byte_perm(x, lc, 0x7604));
            const uint32_t v1 = lds0(__byte_perm(x1, lc, 0x7615));
            const uint32_t v2 = lds1(__byte_perm(x2, lc, 0x7624));
            const uint32_t v3 = lds1(__byte_perm(x3, lc, 0x7635));
            x = v0 ^ v1 ^ v2 ^ v3 ^ rk[4 * r];
        }
        x1 = __shfl_sync(0xffffffffu, x, g0 + ((c + 1u) & 3u));
        x2 = __shfl_sync(0xffffffffu, x, g0 + ((c + 2u) & 3u));
        x3 = __shfl_sync(0xffffffffu, x, g0 + ((c + 3u) & 3u));
        uint32_t o;
        {
            const uint32_t v0 = lds1(__byte_perm(x, lc, 0x7604));   // T2[.]: S in byte 0
            const uint32_t v1 = lds1(__byte_perm(x1, lc, 0x7615));  // T3[.]: S in byte 1
            const uint32_t v2 = lds0(__byte_perm(x2, lc, 0x7624));  // T0[.]: S in byte 2
            const uint32_t v3 = lds0(__byte_perm(x3, lc, 0x7635));  // T1[.]: S in byte 3
            const uint32_t lo = __byte_perm(v0, v1, 0x7650), hi = __byte_perm(v2, v3, 0x7210);
            o = __byte_perm(lo, hi, 0x7610) ^ rk[40] ^ sw;          // (Matyas-Meyer-Oseas feed-forward)
        }
        const uint32_t nxt = __shfl_down_sync(0xffffffffu, o, 1);   // block 1: the next PRG output word (block 2's word 0 after word 3)
        // lanes 0..3: the left child's seed words; lanes 4..7: the right child's; lane 4: tL, lane 8: tR
        const uint32_t child = (k == 0u ? o : __funnelshift_r(o, nxt, 8)) ^ cwc;
        const uint32_t tbit = k == 1u ? ((o & 1u) ^ (t & (uint32_t)(qc.cw_tl >> l) & 1u))
                                      : (((o >> 8) & 1u) ^ (t & (uint32_t)(qc.cw_tr >> l) & 1u));
        sw = __shfl_sync(0xffffffffu, child, (bit ? 4u : 0u) + c);
        t = __shfl_sync(0xffffffffu, tbit, bit ? 8u : 4u);
        if (lane < 4u) reinterpret_cast<uint32_t*>(&path_s[l + 1])[c] = sw;
        if (lane == 0) path_t[l + 1] = (uint8_t)t;
        if (dbg && lane == 0) dbg[14336ull + l] = (unsigned long long)clock64();
    }
    __syncwarp();
}

// ---- top role -------------------------------------------------------------------------------------
// The top of a task (the single-node levels above its node range and the first narrow breadth-first levels)
// is a chain of dependent AES calls that few threads can work on: executed by the tree warps it costs ~30k
// cycles per task during which the lookup pipe idles.  With ap.top_warp set, the last scan warp computes it for
// the tasks ahead while the tree warps are in the depth-first phase of the current one, and hands over the
// nodes of the deepest level that holds at most kHand of them (level Lh) through a double buffer guarded by
// mbarriers.  The tree warps still do task 0 themselves (nothing could hide it).  A 21st warp for this would
// leave 80 registers per thread (a sub-partition holds 16K registers: 6 warps x 32 x 85) and the depth-first
// loop spills; three scan warps plus a dedicated walker cannot keep up with the row stream (both measured).
__device__ __forceinline__ uint32_t nodes_at(uint32_t L, uint64_t a, uint32_t m, uint32_t l) {
    const uint32_t sh = L - l;
    return (uint32_t)(((a + m - 1u) >> sh) - (a >> sh)) + 1u;
}
// level handed over: the deepest level from Ls on with at most kHand nodes (never below the thread roots)
__device__ __forceinline__ uint32_t hand_level(uint32_t L, uint64_t a, uint32_t m, uint32_t Ls) {
    uint32_t Lh = Ls;
    while (Lh < L && nodes_at(L, a, m, Lh + 1u) <= kHand) Lh++;
    return Lh;
}
__device__ __forceinline__ uint32_t hand_full_bar(uint32_t b) { return kAbsTopBars + b * 8u; }
__device__ __forceinline__ uint32_t hand_empty_bar(uint32_t b) { return kAbsTopBars + (2u + b) * 8u; }

struct TaskTop {  // where a task sits in the tree
    uint32_t L, m, Ls, Lh;
    uint64_t a;
    __device__ __forceinline__ TaskTop(const TaskShape& ts, uint64_t row_start, uint32_t top_bits) {
        L = top_bits - ts.D;
        a = row_start >> (ts.D + 1u);
        m = (a + ts.m_nom <= (1ull << L)) ? ts.m_nom : (uint32_t)((1ull << L) - a);
        Ls = single_node_levels(L, a, m);
        Lh = hand_level(L, a, m, Ls);
    }
};

__device__ __forceinline__ bool mbar_test(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    return ok != 0u;
}

// State of the walker (one warp).  Lane 3*j + k computes PRG block k of the j-th node of a pass (coop_block), so
// a level of n nodes takes ceil(n / 10) passes of one AES latency each.
struct TopWalker {
    // (the state lives in shared memory: it is touched between scan units only, and the scan loop needs its registers)
    __device__ __forceinline__ static volatile uint32_t* state() { return reinterpret_cast<volatile uint32_t*>(g_smem2 + kOffTopState); }
    __device__ __forceinline__ static void init() {  // by one thread, before the CTA's first barrier
        volatile uint32_t* st = state();
        st[0] = 1u;  // next task to prepare
        st[1] = 0u;  // Ls of the previous walk
        st[2] = 0u;  // first row of the previous task (64 bit)
        st[3] = 0u;
    }
    __device__ __forceinline__ static bool pending(uint32_t n_tasks) { return state()[0] < n_tasks; }
    // Prepare the next task if its hand-over buffer is free; returns false if there was nothing to do (yet).
    __device__ __forceinline__ static bool try_advance(const QueryConst& qc, uint32_t n_tasks) {
        volatile uint32_t* stt = state();
        const uint32_t it = stt[0];
        if (it >= n_tasks) return false;
        const uint32_t lane = threadIdx.x & 31u, lc = lane_const();
        const uint32_t Ls_prev = stt[1];
        const uint64_t row_prev = ((uint64_t)stt[3] << 32) | stt[2];
        const uint32_t buf = it & 1u, use = (it - 1u) >> 1;
        if (use >= 1u) {  // the tree warps must have taken task it-2's nodes
            uint32_t ok = lane == 0 ? (uint32_t)mbar_test(hand_empty_bar(buf), (use - 1u) & 1u) : 0u;
            ok = __shfl_sync(0xffffffffu, ok, 0);
            if (!ok) return false;
        }
        const uint32_t top_bits = qc.num_bits - 1u;
        const uint32_t dyn_base = smem_addr(g_smem2);
        uint4* path_s = reinterpret_cast<uint4*>(g_smem2 + kOffPath2S);
        uint8_t* path_t = g_smem2 + kOffPath2T;
        uint4* scr_s = reinterpret_cast<uint4*>(g_smem2 + kOffTopScrS);
        uint8_t* scr_t = g_smem2 + kOffTopScrT;
        uint4* hs = reinterpret_cast<uint4*>(g_smem2 + (kAbsHandS - dyn_base)) + buf * kHand;
        uint8_t* ht = g_smem2 + (kAbsHandT - dyn_base) + buf * kHand;
        const TaskIter ti(reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks));
        TaskShape ts;
        uint64_t row_start;
        ti.decode(it, row_start, ts);
        const TaskTop tt(ts, row_start, top_bits);
        const uint32_t k = lane % 3u, jl = lane / 3u;  // block k of the pass's node jl (lanes 30, 31 idle)
        // single-node levels, restarting below the last node shared with the previous walk
        uint32_t cp = 0;
        if (it == 1u) {
            if (lane == 0) {
                path_s[0] = make_uint4(qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]);
                path_t[0] = (uint8_t)qc.t_init;
            }
        } else {
            cp = top_bits + 1u - (64u - (uint32_t)__clzll((long long)(row_start ^ row_prev)));
            cp = cp < Ls_prev ? cp : Ls_prev;
            cp = cp < tt.Ls ? cp : tt.Ls;
        }
        __syncwarp();
        // Levels cp .. Ls-1 hold a single node and live in path_s/path_t (one entry per level, kept for the next
        // restart); from Ls on the levels alternate between the scratch buffer and the hand-over buffer so that
        // level Lh lands in the latter.  Either way a child is kept if it falls into the next level's node range.
        for (uint32_t l = cp; l < tt.Lh; l++) {
            const bool src_hand = ((tt.Lh - l) & 1u) == 0u;
            const uint4* ss = l <= tt.Ls ? &path_s[l] : (src_hand ? hs : scr_s);
            const uint8_t* st = l <= tt.Ls ? &path_t[l] : (src_hand ? ht : scr_t);
            uint4* ds = l < tt.Ls ? &path_s[l + 1] : (src_hand ? scr_s : hs);
            uint8_t* dt = l < tt.Ls ? &path_t[l + 1] : (src_hand ? scr_t : ht);
            const uint32_t sh = tt.L - l;
            const uint64_t lo = tt.a >> sh, lo2 = tt.a >> (sh - 1u);
            const uint32_t n = nodes_at(tt.L, tt.a, tt.m, l), n2 = nodes_at(tt.L, tt.a, tt.m, l + 1u);
            const uint32_t skip = (uint32_t)(lo2 - 2ull * lo);  // 1: the first node's left child is out of range
            for (uint32_t j0 = 0; j0 < n; j0 += 10u) {
                const uint32_t j = j0 + jl;
                if (lane < 30u && j < n) {
                    const uint32_t li = 2u * j - skip, ri = li + 1u;  // (li wraps when the left child is skipped)
                    const bool lv = (j > 0u || skip == 0u) && li < n2, rv = ri < n2;
                    coop_block(qc, k, ss[j], st[j], l, lc, lv ? reinterpret_cast<uint8_t*>(&ds[li]) : nullptr, lv ? &dt[li] : nullptr,
                               rv ? reinterpret_cast<uint8_t*>(&ds[ri]) : nullptr, rv ? &dt[ri] : nullptr);
                }
            }
            __syncwarp();
        }
        if (tt.Lh == tt.Ls && lane == 0) {  // (a task whose thread roots hang below a single node)
            hs[0] = path_s[tt.Ls];
            ht[0] = path_t[tt.Ls];
        }
        if (lane == 0) {
            mbar_arrive(hand_full_bar(buf));  // (release: the warp's stores are visible to the waiters)
            stt[1] = tt.Ls;
            stt[2] = (uint32_t)row_start;
            stt[3] = (uint32_t)(row_start >> 32);
            stt[0] = it + 1u;
        }
        __syncwarp();
        return true;
    }
};

// ---- look-ahead: the top of the NEXT launch's task, by this launch's walker warp (kernels.cuh LaMail / LaTop) -----
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_acquire_gpu_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
constexpr uint32_t kOffLaConst = kOffPath2S;  // the walker's own path arrays are idle while a CTA has a single task
static_assert(sizeof(LaConst) <= 65 * 16 + 80, "LaConst must fit the walker's path arrays");
static_assert(kLaHand == kHand, "hand-over width");

// One warp.  Returns true when there is nothing (more) to do for this launch: the next launch's top has been written,
// or its mail has a different shape / has been overwritten; false if the mail has not arrived yet (poll again later).
// Only called by the walker warp of a CTA with exactly one task: its scratch (walker path arrays, level buffer and the
// hand-over double buffer) is otherwise idle.
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_gpu_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
static_assert(sizeof(LaConst) % 8 == 0, "LaConst is copied in 64-bit words");
constexpr uint64_t kLaNone = 1ull << 63;  // device copy of the mail: "the fetcher found no mail for this sequence number"

// `polls`: how often a CTA other than the fetcher looks for the device copy of the mail before it goes on to scan.
__device__ __noinline__ bool lookahead_next_top(const AnswerPlan& ap, uint32_t top_bits, uint32_t polls) {
    const uint32_t lane = threadIdx.x & 31u, lc = lane_const();
    const uint64_t next = ap.la_seq + 1ull;
    LaConst* la = reinterpret_cast<LaConst*>(g_smem2 + kOffLaConst);
    uint64_t* la64 = reinterpret_cast<uint64_t*>(la);
    LaMail* dmail = ap.la_dmail + (next % kLaTopRing);  // the copy in device memory that CTA 0 publishes
    constexpr uint32_t kWords = sizeof(LaConst) / 8u;
    uint64_t tag = 0, geom = 0;
    // ONE CTA per launch -- the first whose walker gets here -- reads the host-mapped ring: every load is a PCIe round
    // trip, and 148 CTAs doing so cost 0.4 ms per launch (measured).  The claim word only ever grows.
    uint32_t fetcher = 0;
    if (lane == 0) fetcher = atomicMax(reinterpret_cast<unsigned long long*>(&dmail->claim), (unsigned long long)next) < next ? 1u : 0u;
    fetcher = __shfl_sync(0xffffffffu, fetcher, 0);
    if (fetcher) {
        // under a seqlock -- tag, payload, tag again: the host writes 0, payload, tag, and may be re-using the
        // entry for launch next + 64 -- and republishes the entry in device memory for the other CTAs.
        const LaMail* mail = ap.la_mail + (next % kLaMailRing);
        if (lane == 0) {
            tag = ld_acquire_sys_u64(&mail->tag);
            geom = tag == next ? ld_relaxed_sys_u64(&mail->geom) : 0ull;
        }
        tag = __shfl_sync(0xffffffffu, tag, 0);
        geom = __shfl_sync(0xffffffffu, geom, 0);
        if (tag != next) {
            // not there (no query is queued behind this one) or overwritten: tell the other CTAs so that they stop looking
            if (lane == 0) st_release_gpu_u64(&dmail->tag, next | kLaNone);
            return true;
        }
        const uint64_t* src = reinterpret_cast<const uint64_t*>(&mail->c);
        for (uint32_t w = lane; w < kWords; w += 32u) la64[w] = ld_relaxed_sys_u64(src + w);
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        __syncwarp();
        uint64_t tag2 = lane == 0 ? ld_acquire_sys_u64(&mail->tag) : 0ull;
        tag2 = __shfl_sync(0xffffffffu, tag2, 0);
        if (tag2 != next) return true;               // torn read: give up
        // device copy, itself under a seqlock (a later launch's CTA 0 may re-use the entry while a straggler reads it)
        if (lane == 0) st_relaxed_gpu_u64(&dmail->tag, 0ull);
        __threadfence();
        __syncwarp();
        __threadfence();
        uint64_t* dst = reinterpret_cast<uint64_t*>(&dmail->c);
        for (uint32_t w = lane; w < kWords; w += 32u) st_relaxed_gpu_u64(dst + w, la64[w]);
        if (lane == 0) st_relaxed_gpu_u64(&dmail->geom, geom);
        __threadfence();
        __syncwarp();
        if (lane == 0) st_release_gpu_u64(&dmail->tag, next);
        __syncwarp();
    } else {
        for (uint32_t p = 0; p < polls; p++) {
            if (lane == 0) tag = ld_acquire_gpu_u64(&dmail->tag);
            tag = __shfl_sync(0xffffffffu, tag, 0);
            if ((tag & ~kLaNone) >= next) break;
            if (p + 1u < polls) __nanosleep(300);
        }
        if (tag != next) return true;                // the fetcher found nothing, or has not got there in time
        const uint64_t* src = reinterpret_cast<const uint64_t*>(&dmail->c);
        for (uint32_t w = lane; w < kWords; w += 32u) la64[w] = ld_relaxed_gpu_u64(src + w);
        if (lane == 0) geom = ld_relaxed_gpu_u64(&dmail->geom);
        __threadfence();
        __syncwarp();
        uint64_t tag2 = lane == 0 ? ld_acquire_gpu_u64(&dmail->tag) : 0ull;
        tag2 = __shfl_sync(0xffffffffu, tag2, 0);
        geom = __shfl_sync(0xffffffffu, geom, 0);
        if (tag2 != next) return true;               // torn read: give up
    }
    if (geom != ap.la_geom) return true;             // the next launch has another shape: its CTAs walk their own tops
    __syncwarp();
    const uint32_t dyn_base = smem_addr(g_smem2);
    const TaskIter ti(reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks));
    TaskShape ts;
    uint64_t row_start;
    ti.decode(0u, row_start, ts);
    const TaskTop tt(ts, row_start, top_bits);
    if (tt.Lh > kLaMaxLevels) return true;       // (correction words of deeper levels are not in the mail)
    // single-node levels 0 .. Ls: path arrays in the walker's level scratch
    uint4* path_s = reinterpret_cast<uint4*>(g_smem2 + kOffTopScrS);  // [kHand] uint4 >= [65]
    uint8_t* path_t = g_smem2 + kOffTopScrT;                          // [kHand] u8
    if (lane == 0) {
        path_s[0] = make_uint4(la->s_init[0], la->s_init[1], la->s_init[2], la->s_init[3]);
        path_t[0] = (uint8_t)la->t_init;
    }
    __syncwarp();
    quad_path_walk(*la, &la->rk[0][0], lc, tt.a, tt.L, 0u, tt.Ls, path_s, path_t);
    // breadth-first levels Ls .. Lh-1 between the two hand-over buffers (3 lanes per node, 10 nodes per pass)
    uint4* hb_s[2] = {reinterpret_cast<uint4*>(g_smem2 + (kAbsHandS - dyn_base)), reinterpret_cast<uint4*>(g_smem2 + (kAbsHandS - dyn_base)) + kHand};
    uint8_t* hb_t[2] = {g_smem2 + (kAbsHandT - dyn_base), g_smem2 + (kAbsHandT - dyn_base) + kHand};
    const uint32_t k = lane % 3u, jl = lane / 3u;
    for (uint32_t l = tt.Ls; l < tt.Lh; l++) {
        const uint4* ss = l == tt.Ls ? &path_s[l] : hb_s[l & 1u];
        const uint8_t* st = l == tt.Ls ? &path_t[l] : hb_t[l & 1u];
        uint4* ds = hb_s[(l + 1u) & 1u];
        uint8_t* dt = hb_t[(l + 1u) & 1u];
        const uint32_t sh = tt.L - l;
        const uint64_t lo = tt.a >> sh, lo2 = tt.a >> (sh - 1u);
        const uint32_t n = nodes_at(tt.L, tt.a, tt.m, l), n2 = nodes_at(tt.L, tt.a, tt.m, l + 1u);
        const uint32_t skip = (uint32_t)(lo2 - 2ull * lo);
        for (uint32_t j0 = 0; j0 < n; j0 += 10u) {
            const uint32_t j = j0 + jl;
            if (lane < 30u && j < n) {
                const uint32_t li = 2u * j - skip, ri = li + 1u;
                const bool lv = (j > 0u || skip == 0u) && li < n2, rv = ri < n2;
                coop_block(*la, k, ss[j], st[j], l, lc, lv ? reinterpret_cast<uint8_t*>(&ds[li]) : nullptr, lv ? &dt[li] : nullptr,
                           rv ? reinterpret_cast<uint8_t*>(&ds[ri]) : nullptr, rv ? &dt[ri] : nullptr);
            }
        }
        __syncwarp();
    }
    const uint4* fs = tt.Lh == tt.Ls ? &path_s[tt.Ls] : hb_s[tt.Lh & 1u];
    const uint8_t* ft = tt.Lh == tt.Ls ? &path_t[tt.Ls] : hb_t[tt.Lh & 1u];
    const uint32_t n = nodes_at(tt.L, tt.a, tt.m, tt.Lh);
    // The entry is written under a seqlock (ready = 0, nodes, ready = next): with small grids or several streams a launch
    // four further on can be here while a CTA of an earlier one is still copying the entry's previous contents.
    LaTop* out = ap.la_tops + (next % kLaTopRing) * gridDim.x + blockIdx.x;
    if (lane == 0) st_relaxed_gpu_u64(&out->ready, 0ull);
    __threadfence();
    __syncwarp();
    __threadfence();
    for (uint32_t j = lane; j < n; j += 32u) {
        __stcg(&out->seed[j], fs[j]);
        __stcg(&out->t[j], ft[j]);
    }
    if (lane == 0) {
        __stcg(&out->n, n);
        __stcg(&out->lh, tt.Lh);
    }
    __threadfence();
    __syncwarp();
    if (lane == 0) st_release_gpu_u64(&out->ready, next);
    __syncwarp();
    return true;
}

// ---- consumer loop (scan warps all along, tree warps once they have nothing left to expand) ----------
// Work unit = one 32-word group of one slice, claimed in order from a shared counter, so the backlog at
// the end of the kernel is drained by all 20 warps.  The warp that completes a slice's last group hands
// the ring slot back to the tree warps.
template <int NCB>
__device__ __forceinline__ void consume_slices(const QueryConst& qc, const AnswerPlan& ap, uint32_t warp_in_cta, const bool DUAL) {
    const uint32_t lane = threadIdx.x & 31u;
    const ScanLane sl = scan_lane(ap);
    const uint32_t dyn_base = smem_addr(g_smem2);
    uint16_t* list = reinterpret_cast<uint16_t*>(g_smem2 + (kAbsLists - dyn_base)) + warp_in_cta * kListEntries;
    uint32_t* claim = reinterpret_cast<uint32_t*>(g_smem2 + (kAbsClaim - dyn_base));
    uint32_t* done = claim + 1;
    // (only pos, gs0, gs1 stay in registers across scan_group, whose 16 loads in flight need the rest)
    TaskIter ti(reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks));
    uint32_t gs0 = 0, gs1 = 0;  // global slice range of the current task (task ti.pos - 1)
    uint4 acc[NCB];
#pragma unroll
    for (int cb = 0; cb < NCB; cb++) acc[cb] = make_uint4(0, 0, 0, 0);
    long long c_wait = 0;
    const long long c_in = PIRB_AP_DBG(ap) ? clock64() : 0;
    // DUAL: this scan warp also walks the tops of the tasks ahead (TopWalker), before anything else whenever a
    // hand-over buffer is free -- the tree warps will wait for it at the next task boundary, the scan has slack.
    while (true) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(claim, 1u);
        u = __shfl_sync(0xffffffffu, u, 0);
        const uint32_t gsl = u / kGroups, g = u % kGroups;
        bool have = true;
        while (gsl >= gs1) {  // advance to the task that owns this slice
            if (ti.pos >= ti.n) {
                have = false;
                break;
            }
            gs0 = gs1;
            gs1 += 1u << ti.slices_log(ti.pos++);
        }
        if (!have) break;
        const uint32_t slot = gsl & (kRing - 1u), par = (gsl / kRing) & 1u;
        const long long cw0 = (PIRB_AP_DBG(ap) && warp_in_cta == kTreeWarps) ? clock64() : 0;
        if (DUAL) {
            // This warp also walks the tops of the tasks ahead, before anything else whenever a hand-over buffer is
            // free: the tree warps will wait for it at the next task boundary, the scan has slack.  The slice just
            // claimed may belong to a task whose top is still to be walked, so never block on it while tops are pending.
            while (TopWalker::pending(ti.n)) {
                if (TopWalker::try_advance(qc, ti.n)) continue;
                uint32_t ok = lane == 0 ? (uint32_t)mbar_test(full_bar(slot), par) : 0u;
                if (__shfl_sync(0xffffffffu, ok, 0)) break;
            }
        }
        mbar_wait(full_bar(slot), par);
        if (PIRB_AP_DBG(ap) && warp_in_cta == kTreeWarps) c_wait += clock64() - cw0;
        uint64_t grow0;
        uint32_t rows_log, gslog;
        {
            TaskShape ts;
            uint64_t row_start;
            ti.decode(ti.pos - 1u, row_start, ts);
            ts.group_geom(g, row_start, grow0, rows_log, gslog);
        }
        scan_group<(NCB >= 4 ? 8 : 16), NCB>(ap, sl, ring_words(slot), g, (int64_t)grow0 - (int64_t)ap.first_row, rows_log,
                       (gsl - gs0) << gslog, list, acc);
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&done[slot], 1u) == kGroups - 1u) {
                done[slot] = 0u;
                mbar_arrive(empty_bar(slot));
            }
        }
    }
    // (DUAL: no top can be pending here -- units run out only once the tree warps are in the CTA's last task)
    if (PIRB_AP_DBG(ap) && warp_in_cta == kTreeWarps && lane == 0) {  // first scan warp: cycles scanning vs waiting for slices
        PIRB_AP_DBG(ap)[11264ull + blockIdx.x] = ((unsigned long long)(clock64() - c_in - c_wait) << 32) | (unsigned long long)(c_wait & 0xffffffffll);
    }
    fold_acc<NCB>(sl, acc);
}

// ---- tree role ------------------------------------------------------------------------------------
template <int NCB, bool FCW_ODD>
__device__ __forceinline__ void tree_role(const QueryConst& qc, const AnswerPlan& ap, uint32_t lc, long long c_entry) {
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t top_bits = qc.num_bits - 1u;  // rows are (num_bits-1)-bit numbers: the top level always goes left
    const uint64_t first_row = ap.first_row, end_row = ap.first_row + ap.n_rows;
    uint32_t* stk = reinterpret_cast<uint32_t*>(g_smem2 + kOffStack);
    uint4* bfs_s = reinterpret_cast<uint4*>(g_smem2 + kOffBfsS);
    uint8_t* bfs_t = g_smem2 + kOffBfsT;

    uint4* path_s = reinterpret_cast<uint4*>(g_smem2 + kOffPathS);
    uint8_t* path_t = g_smem2 + kOffPathT;

    // (nothing about the task sequence is kept in registers across the depth-first loop but `it`: the
    //  geometry of the current and of the previous task is re-derived from the list in shared memory)
    const TaskIter ti(reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks));
    const uint32_t n_tasks = ti.n;
    uint32_t it = 0;
    uint32_t gs = 0;  // slices published so far (ring position)
    long long c_top = 0, c_bfs = 0, c_dfs = 0, c_tail = 0, c_pub = 0, c_mark = PIRB_AP_DBG(ap) ? clock64() : 0;
    const long long c_begin = c_mark;
#define PHASE(acc)                              \
    if (PIRB_AP_DBG(ap) && tid == 0) {                   \
        const long long now_ = clock64();       \
        acc += now_ - c_mark;                   \
        c_mark = now_;                          \
    }
    for (; it < n_tasks; it++) {
        TaskShape ts;
        uint64_t row_start;
        ti.decode(it, row_start, ts);
        const uint32_t L = top_bits - ts.D;  // tree level of the hi thread roots (2^L nodes, 2^(D+1) rows each)
        const bool hi = warp < ts.hiw;
        const uint32_t D = hi ? ts.D : ts.D - 1u;     // depth of this thread's own subtree (warp-uniform)
        const uint32_t lvl0 = hi ? L : L + 1u;        // tree level of this thread's root
        const uint32_t nlp = 1u << D;                // leaf parents per thread
        const uint32_t rows_log = D + 1u;
        const uint32_t slog = hi ? ts.slog : ts.slog - 1u, lp_log = slog - 1u;
        // the task spans nodes [a, a + m) of level L: hi roots first, then the parents of the lo roots
        const uint64_t a = row_start >> (ts.D + 1u);
        const uint32_t m = (a + ts.m_nom <= (1ull << L)) ? ts.m_nom : (uint32_t)((1ull << L) - a);
        const uint32_t ck = warp & 3u, cj = (warp >> 2) * 32u + lane;  // coop role: block ck of node cj
        const bool coop_warp = ck < 3u && warp < 8u;
        // -- the top, part 1: levels with a single node above the range (the common prefix of a and a+m-1).
        //    They are kept in path_s/path_t, and the walk restarts below the last node shared with the
        //    previous task. --
        const uint32_t Ls = single_node_levels(L, a, m);
        uint32_t l_bfs = Ls;  // first breadth-first level the tree warps expand themselves
        bool la_hit = false;
#ifndef PIRB_NO_LA_CODE
        if (it == 0u && n_tasks == 1u && ap.la_tops != nullptr) {
            // Has the launch before this one left this task's top for us (look-ahead)?  Seqlock read: the sequence number,
            // the nodes (into registers), the sequence number again; anything but two matches and we walk the top ourselves.
            const uint32_t Lh = hand_level(L, a, m, Ls);
            const uint32_t n_h = nodes_at(L, a, m, Lh);
            volatile uint32_t* flag = reinterpret_cast<volatile uint32_t*>(g_smem2 + kOffPathT + 72);
            const LaTop* in = ap.la_tops + (ap.la_seq % kLaTopRing) * gridDim.x + blockIdx.x;
            if (tid == 0) {
                bool ok = ld_acquire_gpu_u64(&in->ready) == ap.la_seq;
                if (ok) ok = __ldcg(&in->lh) == Lh && __ldcg(&in->n) == n_h;
                *flag = ok ? 1u : 0u;
            }
            bar_sync(BAR_TREE, kTreeThreads);
            if (*flag != 0u) {
                uint4 sv = make_uint4(0, 0, 0, 0);
                uint8_t tv = 0;
                if (tid < n_h) {
                    sv = __ldcg(&in->seed[tid]);
                    tv = __ldcg(&in->t[tid]);
                }
                __threadfence();
                bar_sync(BAR_TREE, kTreeThreads);  // (every thread's loads are done; flag has been read by all)
                if (tid == 0) *flag = ld_acquire_gpu_u64(&in->ready) == ap.la_seq ? 1u : 0u;
                bar_sync(BAR_TREE, kTreeThreads);
                la_hit = *flag != 0u;
                if (la_hit) {
                    if (tid < n_h) {
                        bfs_s[(Lh & 1u) * kTreeThreads + tid] = sv;
                        bfs_t[(Lh & 1u) * kTreeThreads + tid] = tv;
                    }
                    l_bfs = Lh;
                }
                bar_sync(BAR_TREE, kTreeThreads);
            }
        }
#endif
        if (la_hit) {
            // (nothing to walk)
        } else if (ap.top_warp && it > 0u) {
            // the top warp has walked this task's top while we were in the previous task's depth-first phase
            const uint32_t Lh = hand_level(L, a, m, Ls);
            const uint32_t buf = it & 1u, use = (it - 1u) >> 1;
            const uint32_t dyn_base = smem_addr(g_smem2);
            mbar_wait(hand_full_bar(buf), use & 1u);
            bar_sync(BAR_TREE, kTreeThreads);  // (the BFS buffers alias the sibling stacks of warps still in the previous task)
            if (tid < nodes_at(L, a, m, Lh)) {
                bfs_s[(Lh & 1u) * kTreeThreads + tid] = reinterpret_cast<const uint4*>(g_smem2 + (kAbsHandS - dyn_base))[buf * kHand + tid];
                bfs_t[(Lh & 1u) * kTreeThreads + tid] = (g_smem2 + (kAbsHandT - dyn_base))[buf * kHand + tid];
            }
            bar_sync(BAR_TREE, kTreeThreads);
            if (tid == 0) mbar_arrive(hand_empty_bar(buf));
            l_bfs = Lh;
        } else {
            uint32_t cp = 0;
            if (it == 0) {
                if (tid == 0) {
                    path_s[0] = make_uint4(qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]);
                    path_t[0] = (uint8_t)qc.t_init;
                }
                bar_sync(BAR_TREE, kTreeThreads);
            } else {
                // common leading bits of the two tasks' first rows (as top_bits+1-bit numbers), capped by how
                // deep the previous and the current single-node walks go
                TaskShape tp;
                uint64_t row_prev;
                ti.decode(it - 1u, row_prev, tp);
                const uint32_t Lp = top_bits - tp.D;
                const uint64_t ap_ = row_prev >> (tp.D + 1u);
                const uint32_t mp = (ap_ + tp.m_nom <= (1ull << Lp)) ? tp.m_nom : (uint32_t)((1ull << Lp) - ap_);
                const uint32_t Ls_prev = single_node_levels(Lp, ap_, mp);
                cp = top_bits + 1u - (64u - (uint32_t)__clzll((long long)(row_start ^ row_prev)));
                cp = cp < Ls_prev ? cp : Ls_prev;
                cp = cp < Ls ? cp : Ls;
            }
            if (warp == 0u) {  // one warp walks the whole chain, the node stays in registers (no CTA barrier per level)
                quad_path_walk(qc, reinterpret_cast<const uint32_t*>(g_smem2 + kOffRk), lc, a, L, cp, Ls, path_s, path_t,
                               (PIRB_AP_DBG(ap) && blockIdx.x == 0 && it == 0) ? PIRB_AP_DBG(ap) : nullptr);
                if (lane == 0) {
                    bfs_s[(Ls & 1u) * kTreeThreads] = path_s[Ls];
                    bfs_t[(Ls & 1u) * kTreeThreads] = path_t[Ls];
                }
            }
            bar_sync(BAR_TREE, kTreeThreads);
        }
        // -- part 2: breadth-first over the node range, level by level.  Level l holds nodes
        //    [a >> (L-l), (a+m-1) >> (L-l)]; a child is kept if it falls into the next level's range. --
        for (uint32_t l = l_bfs; l < L; l++) {
            const uint32_t sh = L - l;
            const uint64_t lo = a >> sh, lo2 = a >> (sh - 1u);
            const uint32_t n = (uint32_t)(((a + m - 1u) >> sh) - lo) + 1u;
            const uint32_t n2 = (uint32_t)(((a + m - 1u) >> (sh - 1u)) - lo2) + 1u;
            const uint32_t skip = (uint32_t)(lo2 - 2ull * lo);  // 1: the first node's left child is out of range
            const uint32_t src = (l & 1u) * kTreeThreads, dst = ((l + 1u) & 1u) * kTreeThreads;
            if (n <= 64u) {
                if (coop_warp && cj < n) {
                    const uint32_t li = 2u * cj - skip, ri = li + 1u;  // (li wraps when the left child is skipped)
                    const bool lv = (cj > 0u || skip == 0u) && li < n2, rv = ri < n2;
                    uint8_t* ls = reinterpret_cast<uint8_t*>(&bfs_s[dst + li]);
                    uint8_t* rs = reinterpret_cast<uint8_t*>(&bfs_s[dst + ri]);
                    coop_block(qc, ck, bfs_s[src + cj], bfs_t[src + cj], l, lc, lv ? ls : nullptr, lv ? &bfs_t[dst + li] : nullptr,
                               rv ? rs : nullptr, rv ? &bfs_t[dst + ri] : nullptr);
                }
            } else if (tid < n) {
                // (one node per thread with the rolled rounds: ~4k cycles per level.  The unrolled AES of the depth-first
                //  loop as an out-of-line copy was measured slower, profiles/r2_shard_probe_variants.txt "widebfs")
                const uint4 sv = bfs_s[src + tid];
                const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
                Children ch;
                expand_node<1>(qc, sw, bfs_t[src + tid], l, lc, ch);
                const uint32_t li = 2u * tid - skip, ri = li + 1u;
                if ((tid > 0u || skip == 0u) && li < n2) {
                    bfs_s[dst + li] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                    bfs_t[dst + li] = (uint8_t)ch.tl;
                }
                if (ri < n2) {
                    bfs_s[dst + ri] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                    bfs_t[dst + ri] = (uint8_t)ch.tr;
                }
            }
            bar_sync(BAR_TREE, kTreeThreads);
        }
        PHASE(c_top)
        // -- the lo roots: one more level below the nodes [m_hi, m) --
        const uint32_t P = (L & 1u) * kTreeThreads, Q = ((L + 1u) & 1u) * kTreeThreads;
        const uint32_t n_par = m > ts.m_hi ? m - ts.m_hi : 0u;
        if (n_par) {
            if (n_par <= 64u) {
                if (coop_warp && cj < n_par) {
                    uint8_t* cs = reinterpret_cast<uint8_t*>(&bfs_s[Q + 2u * cj]);
                    coop_block(qc, ck, bfs_s[P + ts.m_hi + cj], bfs_t[P + ts.m_hi + cj], L, lc, cs, &bfs_t[Q + 2u * cj], cs + 16,
                               &bfs_t[Q + 2u * cj + 1u]);
                }
            } else if (tid < n_par) {
                const uint4 sv = bfs_s[P + ts.m_hi + tid];
                const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
                Children ch;
                expand_node<1>(qc, sw, bfs_t[P + ts.m_hi + tid], L, lc, ch);
                bfs_s[Q + 2u * tid] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                bfs_s[Q + 2u * tid + 1u] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                bfs_t[Q + 2u * tid] = (uint8_t)ch.tl;
                bfs_t[Q + 2u * tid + 1u] = (uint8_t)ch.tr;
            }
            bar_sync(BAR_TREE, kTreeThreads);
        }
        uint32_t s[4] = {0, 0, 0, 0}, t = 0;
        const uint32_t kl = tid - ts.m_hi;  // lo threads: index of the root below the parents
        const bool active = hi ? tid < m : (kl >> 1) < n_par;
        if (active) {
            const uint32_t src = hi ? P + tid : Q + kl;
            const uint4 sv = bfs_s[src];
            s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
            t = bfs_t[src];
        }
        bar_sync(BAR_TREE, kTreeThreads);  // the sibling stack aliases the BFS buffers
        PHASE(c_bfs)

        uint64_t warp_row0;
        {
            uint32_t rl_, sl_;
            ts.group_geom(warp, row_start, warp_row0, rl_, sl_);
        }
        const uint64_t rowbase = warp_row0 + ((uint64_t)lane << rows_log);  // this thread's first row
        const uint64_t warp_row1 = warp_row0 + (32ull << rows_log);
        const bool warp_live = __any_sync(0xffffffffu, active) && warp_row0 < end_row && warp_row1 > first_row;
        const bool inside = row_start >= first_row && row_start + ts.R <= end_row;
        const uint32_t n_slices = 1u << ts.n_slices_log;
        // publish one slice: wait until the scan warps have drained the ring slot, store, arrive
#define PUBLISH(word_)                                                              \
    {                                                                               \
        const uint32_t slot_ = gs & (kRing - 1u), par_ = (gs / kRing) & 1u;         \
        const long long w0_ = (PIRB_AP_DBG(ap) && tid == 0) ? clock64() : 0;                 \
        mbar_wait(empty_bar(slot_), par_ ^ 1u);                                     \
        if (PIRB_AP_DBG(ap) && tid == 0) c_pub += clock64() - w0_;                           \
        ring_words(slot_)[tid] = (word_);                                           \
        __syncwarp();                                                               \
        if (lane == 0) mbar_arrive(full_bar(slot_));                                \
        gs++;                                                                       \
    }
        if (!warp_live) {
            if (NCB != 0)
                for (uint32_t q = 0; q < n_slices; q++) PUBLISH(0u)
        } else {
            uint32_t tst = 0, d = 0, i = 0, bw = 0;
            while (true) {
                const uint32_t lvl = lvl0 + d;
                const uint32_t cw[4] = {qc.cw[lvl][0], qc.cw[lvl][1], qc.cw[lvl][2], qc.cw[lvl][3]};
                const uint32_t cwtl = (uint32_t)(qc.cw_tl >> lvl) & 1u, cwtr = (uint32_t)(qc.cw_tr >> lvl) & 1u;
                uint32_t pair;  // selection bits of the two leaves below a leaf parent
                {
                    uint32_t a[3][4];
                    // FinalCW even: the leaves' control bits do not matter, so a leaf parent skips PRG block 2
                    // (same unrolled code as an inner node, block 2 under a warp-uniform condition)
                    if (FCW_ODD)
                        aes3_main_rounds(qc, s, lc, a);
                    else
                        aes3_main_rounds_cond(qc, s, lc, a, d < D);
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
                    pair = FCW_ODD ? leaf_parent_tail(qc, a, s, t, cw, cwtl, cwtr, lc) : leaf_parent_tail_even(qc, a, s, t, cw, lc);
                }
                bw |= pair << (2u * (i & ((1u << lp_log) - 1u)));
                i++;
                if ((i & ((1u << lp_log) - 1u)) == 0) {  // a slice of every thread's rows is final: publish it
                    const uint32_t q = (i >> lp_log) - 1u;
                    uint32_t m = active ? 0xffffffffu : 0u;
                    if (!inside) m = active ? range_mask32(rowbase + ((uint64_t)q << slog), 1u << slog, first_row, end_row) : 0u;
                    if (NCB != 0) {
                        PUBLISH(bw & m)
                    } else if (bw & m) {
                        // BITS mode: the slice goes straight to the packed bit vector in global memory
                        // (bit i = row first_row + i; the caller zeroed it).  A slice is at most 8 bits.
                        const int64_t bit0 = (int64_t)(rowbase + ((uint64_t)q << slog)) - (int64_t)first_row;
                        const int64_t w0 = bit0 >> 5;  // floor: rows below first_row carry no set bits
                        const uint64_t sh = (uint64_t)(bw & m) << (uint32_t)(bit0 & 31);
                        if ((uint32_t)sh) atomicOr(&ap.bits_out[w0], (uint32_t)sh);
                        if ((uint32_t)(sh >> 32)) atomicOr(&ap.bits_out[w0 + 1], (uint32_t)(sh >> 32));
                    }
                    bw = 0;
                }
                if (i == nlp) break;
                d = D - (uint32_t)(__ffs(i) - 1);  // next subtree root: the sibling stacked at that depth
#pragma unroll
                for (int w = 0; w < 4; w++) s[w] = stk[((d - 1) * 4 + w) * kTreeThreads + tid];
                t = (tst >> (d - 1)) & 1u;
            }
        }
#undef PUBLISH
        PHASE(c_dfs)
    }
    if (PIRB_AP_DBG(ap) && lane == 0) {  // spread of the tree warps' finishing times within the CTA
        const unsigned long long te = (unsigned long long)(clock64() - c_entry);
        atomicMax(PIRB_AP_DBG(ap) + 8192ull + 2ull * blockIdx.x, te);
        atomicMax(PIRB_AP_DBG(ap) + 8192ull + 2ull * blockIdx.x + 1ull, (1ull << 40) - te);
        // scan backlog at that moment: units published minus units claimed
        const uint32_t claimed = *reinterpret_cast<volatile uint32_t*>(g_smem2 + (kAbsClaim - smem_addr(g_smem2)));
        const uint32_t total_u = gs * kGroups;
        atomicMax(PIRB_AP_DBG(ap) + 10240ull + blockIdx.x, (unsigned long long)(total_u > claimed ? total_u - claimed : 0u) + 1ull);
    }
    // nothing left to expand: help the scan warps drain the ring
    if constexpr (NCB != 0) consume_slices<NCB>(qc, ap, warp, false);
    PHASE(c_tail)
#undef PHASE
    if (PIRB_AP_DBG(ap) && tid == 0) {
        unsigned long long* o = PIRB_AP_DBG(ap) + 8ull * blockIdx.x;
        o[0] = (unsigned long long)c_top; o[1] = (unsigned long long)c_bfs; o[2] = (unsigned long long)c_pub;
        o[3] = (unsigned long long)c_dfs; o[4] = (unsigned long long)c_tail; o[5] = (unsigned long long)(clock64() - c_begin);
        o[6] = it; o[7] = (unsigned long long)(c_begin - c_entry);
    }
}

template <int NCB, bool FCW_ODD>
__global__ void __launch_bounds__(kAnswerThreads, 1)
answer_kernel(const __grid_constant__ QueryConst qc, const __grid_constant__ AnswerPlan ap,
              const uint32_t* __restrict__ t0) {
    const uint32_t base = smem_addr(g_smem2);
    // Programmatic dependent launch: the next answer kernel in the stream may start its CTAs on SMs this launch has
    // already left (table fill, tree top and expansion read only the database and the key).  Everything a launch
    // WRITES to global memory comes after griddepcontrol.wait, i.e. after the launch before it has completed, so
    // the accumulators and counters are used strictly in stream order.
    asm volatile("griddepcontrol.launch_dependents;");
    if constexpr (NCB == 0) asm volatile("griddepcontrol.wait;" ::: "memory");  // (BITS mode writes as it goes)
    const long long c_entry = PIRB_AP_DBG(ap) ? clock64() : 0;
    if (PIRB_AP_DBG(ap) && threadIdx.x == 0) {  // CTA start on the GPU-wide timer (ns): launch stagger across SMs
        unsigned long long gt;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        PIRB_AP_DBG(ap)[12288ull + blockIdx.x] = gt;
    }
    {
        uint32_t dyn;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
        if (base + dyn < kSmemEnd || base + kScratchEnd > kTableBase) __trap();
    }
    fill_tables(t0);
    uint32_t* red = red_words();
    const uint32_t words_out = (uint32_t)(ap.row_bytes >> 2);
    if constexpr (NCB != 0)
        for (uint32_t i = threadIdx.x; i < words_out; i += kAnswerThreads) red[i] = 0;
    if (threadIdx.x < 3 * 44) reinterpret_cast<uint32_t*>(g_smem2 + kOffRk)[threadIdx.x] = qc.rk[threadIdx.x / 44][threadIdx.x % 44];
    if (threadIdx.x == kAnswerThreads - 1) build_task_list(ap, reinterpret_cast<uint32_t*>(g_smem2 + kOffTasks));
    if (threadIdx.x == kAnswerThreads - 2) {
        for (uint32_t k = 0; k < kRing; k++) {
            mbar_init(full_bar(k), kTreeWarps);
            mbar_init(empty_bar(k), 1);
        }
        for (uint32_t k = 0; k < 4; k++) mbar_init(kAbsTopBars + k * 8u, 1);  // hand_full[2], hand_empty[2]
        TopWalker::init();
        uint32_t* claim = reinterpret_cast<uint32_t*>(g_smem2 + (kAbsClaim - base));
        for (uint32_t k = 0; k < 1 + kRing; k++) claim[k] = 0;
    }
    __syncthreads();
    if (threadIdx.x < kTreeThreads) {
        tree_role<NCB, FCW_ODD>(qc, ap, lane_const(), c_entry);
    } else {
        // the last scan warp doubles as the top walker
        const bool dual = ap.top_warp != 0u && threadIdx.x >= kAnswerThreads - 32;
        if constexpr (NCB != 0) {
            // Look-ahead (the walker warp of a CTA with one task, ap.la_mail set): until the first slice arrives the scan
            // warps idle anyway, so that is when the next launch's top is walked.  One attempt per launch: under load the
            // next launch's mail has been waiting since long before this one started.
#ifndef PIRB_NO_LA_CODE
            if (dual && ap.la_mail != nullptr && *reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks) == 1u)
                lookahead_next_top(ap, qc.num_bits - 1u, 16u);
#endif
            consume_slices<NCB>(qc, ap, threadIdx.x >> 5, dual);
        } else if (dual) {
            const uint32_t n_tasks = *reinterpret_cast<const uint32_t*>(g_smem2 + kOffTasks);
            while (TopWalker::pending(n_tasks)) TopWalker::try_advance(qc, n_tasks);
        }
    }
    if constexpr (NCB != 0) {  // (BITS mode has nothing to fold)
        __syncthreads();
        asm volatile("griddepcontrol.wait;" ::: "memory");
        if (ap.arrivals && !ap.final_out) {  // (the slot may live on a peer GPU and several GPUs XOR into it: system scope)
            for (uint32_t i = threadIdx.x; i < words_out; i += kAnswerThreads) {
                const uint32_t v = red[i];
                if (v) asm volatile("red.relaxed.sys.global.xor.b32 [%0], %1;" ::"l"(ap.out + i), "r"(v) : "memory");
            }
        } else {
            for (uint32_t i = threadIdx.x; i < words_out; i += kAnswerThreads) {
                const uint32_t v = red[i];
                if (v) atomicXor(&ap.out[i], v);
            }
        }
        if (ap.final_out) {
            // Single-GPU call: every CTA's XORs are fenced and counted; the last CTA reads the finished words
            // (atomicExch: the read also re-zeroes the accumulator for the next query) and stores them to the
            // caller's buffer -- device memory, or a mapped pinned host buffer that the host sees once the kernel
            // has completed.  One launch per query: no memset before and no copy after the kernel.
            __threadfence();
            __syncthreads();
            uint32_t* flag = reinterpret_cast<uint32_t*>(g_smem2 + (kAbsClaim - smem_addr(g_smem2)));  // (free by now)
            if (threadIdx.x == 0) {
                const bool last = atomicAdd(ap.cta_done, 1u) == gridDim.x - 1u;
                if (last) *ap.cta_done = 0u;  // ready for the launch that reuses this counter
                *flag = last ? 1u : 0u;
            }
            __syncthreads();
            if (*flag && !ap.arrivals) {
                __threadfence();
                for (uint32_t i = threadIdx.x; i < words_out; i += kAnswerThreads) ap.final_out[i] = atomicExch(&ap.out[i], 0u);
                __threadfence_system();
            } else if (*flag) {
                // Multi-GPU combine, folded: final_out is the combine slot (possibly on a peer GPU) that all shards XOR
                // into.  Only this CTA pays the round trip: forward the shard's sum, then one thread fences with system
                // scope (cumulative over the barrier) and publishes the arrival.
                __threadfence();
                for (uint32_t i = threadIdx.x; i < words_out; i += kAnswerThreads) {
                    const uint32_t v = atomicExch(&ap.out[i], 0u);
                    if (v) asm volatile("red.relaxed.sys.global.xor.b32 [%0], %1;" ::"l"(ap.final_out + i), "r"(v) : "memory");
                }
                __syncthreads();
                if (threadIdx.x == 0) {
                    asm volatile("fence.acq_rel.sys;" ::: "memory");
                    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(ap.arrivals) : "memory");
                }
            }
        } else if (ap.arrivals) {
            // Multi-GPU combine: this shard's partial is complete once every CTA's XORs are visible.  The CTA's
            // threads meet at a barrier and ONE thread fences with system scope (the fence is cumulative over what
            // the barrier ordered before it: the pattern of a cooperative-groups grid sync and of NCCL's flag stores;
            // round 1 had all 640 threads execute the system fence), counts the CTA in with release/acquire at GPU
            // scope, and the last CTA of the launch publishes the arrival with a system-scope release
            // (kernels.cu wait kernels acquire it on the owner).
            // (Folding the CTAs' partials in local memory first and forwarding one sum per shard was measured on
            // 8 GPUs: no difference, the 148 x 64 remote atomics per shard are not a bottleneck.)
            __syncthreads();
            if (threadIdx.x == 0) {
                uint32_t old;
                asm volatile("fence.acq_rel.sys;" ::: "memory");
                asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(ap.cta_done) : "memory");
                if (old == gridDim.x - 1u) {
                    *ap.cta_done = 0u;  // ready for the launch that reuses this counter
                    asm volatile("fence.acq_rel.sys;" ::: "memory");
                    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(ap.arrivals) : "memory");
                }
            }
        }
    }
    if (PIRB_AP_DBG(ap)) {
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long gt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
            PIRB_AP_DBG(ap)[13312ull + blockIdx.x] = gt;
        }
    }
}

bool answer_kernel_supports(uint64_t row_bytes, uint32_t D) {
    if (row_bytes < 16 || row_bytes > kMaxRowBytes || (row_bytes & 15)) return false;  // any multiple of 16 up to 4 KB
    return D <= (uint32_t)kV2MaxDepth;
}

template <int NCB>
static cudaError_t configure_answer_ncb() {
    cudaError_t e = cudaFuncSetAttribute(answer_kernel<NCB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemEnd);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(answer_kernel<NCB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemEnd);
    return e;
}

cudaError_t configure_answer_kernel() {
    cudaError_t e = configure_answer_ncb<0>();
    if (e == cudaSuccess) e = configure_answer_ncb<1>();
    if (e == cudaSuccess) e = configure_answer_ncb<2>();
    if (e == cudaSuccess) e = configure_answer_ncb<4>();
    if (e == cudaSuccess) e = configure_answer_ncb<8>();
    return e;
}

uint64_t answer_tasks(uint64_t first_row, uint64_t n_rows, uint32_t D, uint32_t hi_warps, uint32_t m_task) {
    return answer_tasks_of(first_row, n_rows, D, hi_warps, m_task);
}
uint32_t answer_task_nodes(uint32_t hi_warps, uint32_t m_task) { return task_nodes(hi_warps, m_task); }
uint64_t answer_chunks(uint64_t first_row, uint64_t n_rows, uint32_t Dmin) {
    const uint32_t cl = 10u + Dmin;
    return ((first_row + n_rows + ((1ull << cl) - 1ull)) >> cl) - (first_row >> cl);
}

cudaError_t launch_answer(const QueryConst& qc, const AnswerPlan& ap, const uint32_t* d_t0, uint32_t smem_base,
                          int num_sms, cudaStream_t st, bool overlap) {
    if (ap.n_rows == 0) return cudaSuccess;
    if (ap.D > (uint32_t)kV2MaxDepth || ap.D + 1u > qc.num_bits || ap.hi_warps > (uint32_t)kTreeWarps || (ap.hi_warps & 3u) ||
        (ap.hi_warps != 0u && ap.hi_warps < (uint32_t)kTreeWarps && ap.D == 0u) || (ap.hi_warps == 0u && ap.Dmin > ap.D) ||
        ap.slog_cap < 2u || ap.slog_cap > kSliceLog)  // (2: the lo warps of a mixed task publish half-size slices of >= 2 rows)
        return cudaErrorInvalidValue;
    // units the CTAs share out: tasks (uniform split) or chunks (binary split).  A CTA's task list holds
    // kMaxTasks entries: very large shards run more CTAs than SMs (in waves) rather than overflow it.
    const uint64_t units = ap.hi_warps ? answer_tasks(ap.first_row, ap.n_rows, ap.D, ap.hi_warps, ap.m_task)
                                       : answer_chunks(ap.first_row, ap.n_rows, ap.Dmin);
    uint64_t g = units < (uint64_t)num_sms ? units : (uint64_t)num_sms;
    const uint64_t gmin = (units + (kMaxTasks - 1)) / kMaxTasks;
    if (g < gmin) g = gmin;
    const size_t smem = kSmemEnd - smem_base;
    // two code paths per row geometry: keys with an odd FinalCW need the leaves' control bits, keys with an
    // even one do not (aes_unrolled.cuh, leaf_parent_tail_even)
    const bool odd = qc.fcw_parity != 0 || (PIRB_AP_PAD(ap) & 2u);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)g);
    cfg.blockDim = dim3(kAnswerThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap ? 1 : 0;
    cudaError_t lerr = cudaSuccess;
#define PIRB_LAUNCH(NCB)                                                                           \
    do {                                                                                           \
        if (odd)                                                                                   \
            lerr = cudaLaunchKernelEx(&cfg, answer_kernel<NCB, true>, qc, ap, d_t0);               \
        else                                                                                       \
            lerr = cudaLaunchKernelEx(&cfg, answer_kernel<NCB, false>, qc, ap, d_t0);              \
    } while (0)
    if (ap.bits_out)
        PIRB_LAUNCH(0);
    else if (ap.row_bytes <= 512)
        PIRB_LAUNCH(1);
    else if (ap.row_bytes <= 1024)
        PIRB_LAUNCH(2);
    else if (ap.row_bytes <= 2048)
        PIRB_LAUNCH(4);
    else
        PIRB_LAUNCH(8);
#undef PIRB_LAUNCH
    count_launch();
    return lerr != cudaSuccess ? lerr : cudaGetLastError();
}

}  // namespace pirb
