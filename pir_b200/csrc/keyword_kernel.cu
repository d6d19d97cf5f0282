// keyword_kernel.cu — DPF evaluation at arbitrary points (keyword queries, db.go:132-135: numBits = 32 and
// the point is Keywords[row]; also the batched dpf API) with prefix sharing between consecutive points, and, for the
// answer path, the row XOR of db.go:91-104 fused into the same kernel.
//
// The reference walks all numBits levels for every row.  Rows of a keyword database are (usually) sorted by
// keyword, so consecutive points share a long prefix of their root-to-leaf paths.  Every thread takes a run of
// consecutive points and walks them like a depth-first traversal of the sparse tree they span: whenever it turns
// LEFT it keeps the right child of that node (a sibling stack: the deepest kKwSmemLevels levels in shared memory,
// one conflict-free column per thread, the rarer shallower ones in thread-private memory); the walk for the next
// point restarts from the stacked sibling at the first level where the two points differ: numBits - 1 - cp node
// expansions instead of numBits, where cp = number of common leading bits (about 4.5 instead of 32 for 2^28 sorted
// keywords in a 32-bit domain).  A point that does not continue the traversal (smaller than its predecessor, or
// the sibling it needs was never stacked) restarts at the root: results are identical for ANY order of points, only
// the speed depends on it.  Each lane runs its own (point, level) state machine, one node expansion per step, so
// lanes with short and long restarts do not wait for each other.
//
// Leaf parents are not expanded in that walk.  The deepest node of every point needs ONE child, and of it only 8 seed
// bytes and (odd FinalCW, or full values wanted) the control bit: 150-300 table lookups instead of 458 -- but a lookup
// instruction costs a shared-memory wavefront however few lanes execute it, and the lanes of a warp reach their leaf
// parents at different steps and want different sides.  So a lane that arrives at a leaf parent puts (seed, t, point
// index) into its warp's shared-memory queue for that side and goes straight on to its next point; whenever a queue
// holds 32 entries the whole warp computes them with uniform code (aes_unrolled.cuh leaf_child_left / _right).
//
// FUSED mode (the keyword answer path): when a lane reaches a leaf whose selection bit is set, the warp fetches that
// row cooperatively (lane = 16-byte chunk, 32/c rows per load for rows of c chunks) and XORs it into per-lane
// accumulators: no byte-per-row bit vector goes through HBM and no second kernel runs.
#include "kernels.cuh"
#include "aes_unrolled.cuh"

namespace pirb {

extern __shared__ __align__(16) uint8_t g_kwsmem[];

constexpr int kKwThreads = 512;
constexpr int kKwWarps = kKwThreads / 32;
constexpr uint32_t kKwLevelBytes = kKwThreads * 16;              // one stacked sibling per thread per level
constexpr uint32_t kKwSmemLevels = 5;                            // sibling levels below the tables: [base, base + 40K)
constexpr uint32_t kKwQueueEntries = 64;                         // per warp per side: < 32 before a step, <= 63 after it
constexpr uint32_t kKwQueueBytes = kKwQueueEntries * 20;         // [64] uint4 seeds + [64] u32 (t << 31 | point index)
constexpr uint32_t kKwOffQueue1 = kKwSmemLevels * kKwLevelBytes;               // right-child queues: below the tables, behind the stacks
constexpr uint32_t kKwAbsQueue0 = kTableEnd;                                    // left-child queues: above the tables
constexpr uint32_t kKwAbsCw = kKwAbsQueue0 + kKwWarps * kKwQueueBytes;         // [64][4] u32 correction words
constexpr uint32_t kKwAbsRed = kKwAbsCw + kMaxLevels * 16;                      // [128] u32: the CTA's partial answer
constexpr uint32_t kKwSmemEnd = kKwAbsRed + 512;
static_assert(kKwSmemEnd <= 227u * 1024u, "keyword kernel shared memory");
static_assert(2048u + kKwOffQueue1 + kKwWarps * kKwQueueBytes <= kTableBase, "keyword kernel scratch below the tables (the kernel checks its real base)");

// shared-memory address of word w of thread tid's sibling at stack level sl (conflict-free: tid is the fast index)
__device__ __forceinline__ uint32_t kw_sib_addr(uint32_t dyn_base, uint32_t sl, uint32_t w, uint32_t tid) {
    return dyn_base + sl * kKwLevelBytes + (w * kKwThreads + tid) * 4u;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}

// Everything a batch of queued leaf parents needs besides the queue itself.
struct KwOut {
    int64_t* values;
    uint8_t* bits;
    const uint8_t* db;
    uint32_t row_bytes, c, rpl, my_slot, my_chunk;  // FUSED: lane = (row slot, 16-byte chunk)
};

// The first `nvalid` entries of the calling warp's queue `qbase` (from ring position `head`): one child (side) of each leaf
// parent, its selection bit / value written out, and in FUSED mode the selected rows XORed into acc.
template <bool FUSED>
__device__ __forceinline__ void kw_leaf_batch(const QueryConst& qc, uint32_t qbase, uint32_t head, uint32_t nvalid, uint32_t side,
                                              bool need_t, uint32_t lc, const KwOut& o, uint4& acc) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t nb = qc.num_bits, lvl = nb - 1u;
    const bool valid = lane < nvalid;
    const uint32_t e = (head + lane) & (kKwQueueEntries - 1u);
    uint4 sv = make_uint4(0, 0, 0, 0);
    uint32_t meta = 0;
    if (valid) {
        sv = lds128(qbase + e * 16u);
        meta = lds32(qbase + kKwQueueEntries * 16u + e * 4u);
    }
    const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
    const uint32_t t = meta >> 31;
    const uint64_t i = meta & 0x7fffffffu;
    const uint32_t cw0 = lds32(kKwAbsCw + (lvl * 4u + 0u) * 4u), cw1 = lds32(kKwAbsCw + (lvl * 4u + 1u) * 4u);
    uint32_t w0, w1, tc;
    if (side == 0u)
        leaf_child_left(qc, s, t, cw0, cw1, (uint32_t)(qc.cw_tl >> lvl) & 1u, need_t, lc, w0, w1, tc);
    else
        leaf_child_right(qc, s, t, cw0, cw1, (uint32_t)(qc.cw_tr >> lvl) & 1u, need_t, lc, w0, w1, tc);
    uint32_t sel = 0;
    if (valid) {
        if (o.values) o.values[i] = leaf_value(w0, w1, tc, qc.final_cw, qc.share_number);
        sel = leaf_select_bit(w0, w1, tc, qc.fcw_parity);  // (tc is only used when FinalCW is odd: need_t covers that)
        if (o.bits) o.bits[i] = (uint8_t)sel;
    }
    if (FUSED) {
        // rows selected in this batch, fetched by the whole warp: slot k of an iteration takes the k-th of them
        const uint32_t bal = __ballot_sync(0xffffffffu, valid && sel);
        const uint32_t cnt = __popc(bal);
        for (uint32_t k0 = 0; k0 < cnt; k0 += o.rpl) {
            const uint32_t k = k0 + o.my_slot;
            const uint32_t src = k < cnt ? __fns(bal, 0u, (int)k + 1) : 0u;
            const uint32_t row = __shfl_sync(0xffffffffu, (uint32_t)i, src & 31u);
            if (k < cnt && o.my_slot < o.rpl) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(o.db + (uint64_t)row * o.row_bytes) + o.my_chunk);
                acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
            }
        }
    }
}

template <bool FUSED>
__global__ void __launch_bounds__(kKwThreads, 1)
keyword_walk_kernel(const __grid_constant__ QueryConst qc, const uint64_t* __restrict__ xs, uint64_t x_first, uint64_t n,
                    uint32_t run_len, int64_t* __restrict__ values, uint8_t* __restrict__ bits,
                    const uint32_t* __restrict__ t0, const uint8_t* __restrict__ db, uint32_t row_bytes,
                    uint32_t* __restrict__ out) {
    const uint32_t dyn_base = smem_addr(g_kwsmem);
    {
        uint32_t dyn;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
        if (dyn_base + kKwOffQueue1 + kKwWarps * kKwQueueBytes > kTableBase || dyn_base + dyn < kKwSmemEnd) __trap();
    }
    fill_tables(t0);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t nb = qc.num_bits;
    // correction words in shared memory: the lanes of a warp sit at different levels, and an indexed constant-bank
    // load serialises over distinct addresses
    for (uint32_t w = tid; w < nb * 4u; w += kKwThreads) sts32(kKwAbsCw + w * 4u, qc.cw[w >> 2][w & 3u]);
    if (FUSED)
        for (uint32_t w = tid; w < 128u; w += kKwThreads) sts32(kKwAbsRed + w * 4u, 0u);
    __syncthreads();
    const uint32_t lc = lane_const();
    const uint64_t dom_mask = nb >= 64u ? ~0ull : ((1ull << nb) - 1ull);
    const uint32_t leaf_lvl = nb - 1u;  // level of the leaf parents: never expanded in the walk, queued instead
    const uint32_t smem_l0 = leaf_lvl > kKwSmemLevels ? leaf_lvl - kKwSmemLevels : 0u;  // levels [smem_l0, leaf_lvl): siblings in shared memory
    const bool need_t = qc.fcw_parity != 0u || values != nullptr;                        // (see leaf_child_left)
    uint4 deep[kMaxLevels];  // siblings of the levels above smem_l0 (thread-private, rarely touched)
    KwOut o;
    o.values = values; o.bits = bits; o.db = db; o.row_bytes = row_bytes;
    o.c = FUSED ? row_bytes >> 4 : 1u;
    o.rpl = FUSED ? 32u / o.c : 1u;
    o.my_slot = lane / o.c;
    o.my_chunk = lane - o.my_slot * o.c;
    uint4 acc = make_uint4(0, 0, 0, 0);
    // this warp's two queues (ring position of the oldest entry, entries); warp-uniform, kept by every lane
    const uint32_t qbase[2] = {kKwAbsQueue0 + warp * kKwQueueBytes, dyn_base + kKwOffQueue1 + warp * kKwQueueBytes};
    uint32_t qhead[2] = {0u, 0u}, qcnt[2] = {0u, 0u};
    const uint64_t n_runs = (n + run_len - 1) / run_len;
    // the run loop and the step loop are warp-uniform (vote on a per-lane `active` flag): the lanes of a warp
    // re-converge every step, so a step always expands one node for every lane that still has work
    for (uint64_t run0 = (uint64_t)blockIdx.x * kKwThreads + (tid & ~31u); run0 < n_runs; run0 += (uint64_t)gridDim.x * kKwThreads) {
        const uint64_t run = run0 + lane;
        bool active = run < n_runs;
        uint64_t i = active ? run * run_len : 0;
        const uint64_t i_end = i + run_len < n ? i + run_len : n;
        uint64_t x = active ? ((xs ? xs[i] : x_first + i) & dom_mask) : 0;
        uint32_t s[4] = {qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]};
        uint32_t t = qc.t_init, lvl = 0;
        uint64_t vmask = 0, tmask = 0;  // bit l: the right child of the path node at level l is stacked / its control bit
        while (__any_sync(0xffffffffu, active)) {
            if (active && lvl < leaf_lvl) {
                uint32_t cw[4];
#pragma unroll
                for (int w = 0; w < 4; w++) cw[w] = lds32(kKwAbsCw + (lvl * 4u + (uint32_t)w) * 4u);
                const uint32_t cwtl = (uint32_t)(qc.cw_tl >> lvl) & 1u, cwtr = (uint32_t)(qc.cw_tr >> lvl) & 1u;
                uint32_t a[3][4];
                aes3_main_rounds(qc, s, lc, a);
                Children ch;
                inner_tail(qc, a, s, t, cw, cwtl, cwtr, lc, ch);
                const uint32_t bit = (uint32_t)(x >> (nb - 1u - lvl)) & 1u;  // getBit, dpf/common.go:56-58
                if (bit == 0u) {
                    // turning left: stack the right child for the points that come later in the traversal
                    if (lvl >= smem_l0) {
                        const uint32_t sl = lvl - smem_l0;
#pragma unroll
                        for (int w = 0; w < 4; w++) sts32(kw_sib_addr(dyn_base, sl, (uint32_t)w, tid), ch.sr[w]);
                    } else {
                        deep[lvl] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                    }
                    vmask |= 1ull << lvl;
                    tmask = (tmask & ~(1ull << lvl)) | ((uint64_t)ch.tr << lvl);
#pragma unroll
                    for (int w = 0; w < 4; w++) s[w] = ch.sl[w];
                    t = ch.tl;
                } else {
                    vmask &= ~(1ull << lvl);
#pragma unroll
                    for (int w = 0; w < 4; w++) s[w] = ch.sr[w];
                    t = ch.tr;
                }
                lvl++;
            }
            // lanes that hold a leaf parent queue it for the side their point is on ...
            const bool atleaf = active && lvl == leaf_lvl;
            const uint32_t my_side = (uint32_t)x & 1u;
#pragma unroll
            for (uint32_t side = 0; side < 2u; side++) {
                const bool mine = atleaf && my_side == side;
                const uint32_t bal = __ballot_sync(0xffffffffu, mine);
                if (mine) {
                    const uint32_t e = (qhead[side] + qcnt[side] + __popc(bal & ((1u << lane) - 1u))) & (kKwQueueEntries - 1u);
                    sts128(qbase[side] + e * 16u, make_uint4(s[0], s[1], s[2], s[3]));
                    sts32(qbase[side] + kKwQueueEntries * 16u + e * 4u, (t << 31) | (uint32_t)i);
                }
                qcnt[side] += __popc(bal);
            }
            // ... and go straight on to their next point: restart from the sibling stacked where the two points part
            if (atleaf) {
                i++;
                if (i >= i_end) {
                    active = false;
                } else {
                    const uint64_t xn = (xs ? xs[i] : x_first + i) & dom_mask;
                    const uint64_t diff = x ^ xn;
                    x = xn;
                    const uint32_t cp = diff ? nb - (64u - (uint32_t)__clzll((long long)diff)) : nb;  // common leading bits
                    if (cp < leaf_lvl) {  // (otherwise: the same leaf parent again -- its other child or the same one)
                        const uint32_t need = (uint32_t)(xn >> (nb - 1u - cp)) & 1u;
                        if (need == 1u && ((vmask >> cp) & 1ull)) {
                            if (cp >= smem_l0) {
                                const uint32_t sl = cp - smem_l0;
#pragma unroll
                                for (int w = 0; w < 4; w++) s[w] = lds32(kw_sib_addr(dyn_base, sl, (uint32_t)w, tid));
                            } else {
                                const uint4 sv = deep[cp];
                                s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
                            }
                            t = (uint32_t)(tmask >> cp) & 1u;
                            lvl = cp + 1u;
                            vmask &= (1ull << cp) - 1ull;  // level cp is now a right turn; deeper entries belong to the old path
                        } else {
                            // not a continuation of the traversal (unsorted input): walk from the root again
                            s[0] = qc.s_init[0]; s[1] = qc.s_init[1]; s[2] = qc.s_init[2]; s[3] = qc.s_init[3];
                            t = qc.t_init;
                            lvl = 0;
                            vmask = 0;
                        }
                    }
                }
            }
            __syncwarp();
#pragma unroll
            for (uint32_t side = 0; side < 2u; side++) {
                if (qcnt[side] >= 32u) {  // (warp-uniform)
                    kw_leaf_batch<FUSED>(qc, qbase[side], qhead[side], 32u, side, need_t, lc, o, acc);
                    qhead[side] = (qhead[side] + 32u) & (kKwQueueEntries - 1u);
                    qcnt[side] -= 32u;
                    __syncwarp();
                }
            }
        }
    }
    // what is left in the queues (fewer than 32 entries per side, twice if a step's arrivals pushed one past 32)
#pragma unroll
    for (uint32_t side = 0; side < 2u; side++) {
        while (qcnt[side] != 0u) {
            const uint32_t nv = qcnt[side] < 32u ? qcnt[side] : 32u;
            kw_leaf_batch<FUSED>(qc, qbase[side], qhead[side], nv, side, need_t, lc, o, acc);
            qhead[side] = (qhead[side] + nv) & (kKwQueueEntries - 1u);
            qcnt[side] -= nv;
            __syncwarp();
        }
    }
    if (FUSED) {
        if (o.my_slot < o.rpl && (acc.x | acc.y | acc.z | acc.w)) {
            const uint32_t ra = kKwAbsRed + o.my_chunk * 16u;
            asm volatile("red.shared.xor.b32 [%0], %1;" ::"r"(ra), "r"(acc.x) : "memory");
            asm volatile("red.shared.xor.b32 [%0], %1;" ::"r"(ra + 4u), "r"(acc.y) : "memory");
            asm volatile("red.shared.xor.b32 [%0], %1;" ::"r"(ra + 8u), "r"(acc.z) : "memory");
            asm volatile("red.shared.xor.b32 [%0], %1;" ::"r"(ra + 12u), "r"(acc.w) : "memory");
        }
        __syncthreads();
        for (uint32_t w = tid; w < (row_bytes >> 2); w += kKwThreads) {
            const uint32_t v = lds32(kKwAbsRed + w * 4u);
            if (v) atomicXor(&out[w], v);
        }
    }
}

cudaError_t configure_keyword_kernel() {
    cudaError_t e = cudaFuncSetAttribute(keyword_walk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKwSmemEnd);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(keyword_walk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKwSmemEnd);
    return e;
}

bool keyword_fused_supports(uint64_t row_bytes) { return row_bytes >= 16 && row_bytes <= 512 && (row_bytes & 15) == 0; }

// values and/or bits for n points; xs == nullptr means x = x_first + i.  With db != nullptr (row_bytes a multiple of
// 16 up to 512) the selected rows db[i * row_bytes ..] are XORed into out (zeroed or accumulating, the caller's choice).
cudaError_t launch_keyword_walk(const QueryConst& qc, const uint64_t* xs, uint64_t x_first, uint64_t n, int64_t* values,
                                uint8_t* bits, const uint32_t* d_t0, uint32_t smem_base, int num_sms, cudaStream_t st,
                                const uint8_t* db, uint64_t row_bytes, uint32_t* out) {
    if (n == 0) return cudaSuccess;
    if (db && !keyword_fused_supports(row_bytes)) return cudaErrorInvalidValue;
    if (n >= (1ull << 31)) return cudaErrorInvalidValue;  // (a queue entry keeps the point index in 31 bits)
    // one run per thread: the points are split evenly over all resident threads, so every SM gets the same
    // number of points and the first (full) walk of a run is amortised over as many points as possible
    const uint64_t threads = (uint64_t)num_sms * kKwThreads;
    uint64_t run_len = (n + threads - 1) / threads;
    if (run_len < 1) run_len = 1;
    const uint64_t n_runs = (n + run_len - 1) / run_len;
    uint64_t blocks = (n_runs + kKwThreads - 1) / kKwThreads;
    if (blocks > (uint64_t)num_sms) blocks = (uint64_t)num_sms;
    const size_t smem = kKwSmemEnd - smem_base;
    if (db)
        keyword_walk_kernel<true><<<(unsigned)blocks, kKwThreads, smem, st>>>(qc, xs, x_first, n, (uint32_t)run_len, values, bits, d_t0,
                                                                              db, (uint32_t)row_bytes, out);
    else
        keyword_walk_kernel<false><<<(unsigned)blocks, kKwThreads, smem, st>>>(qc, xs, x_first, n, (uint32_t)run_len, values, bits, d_t0,
                                                                               nullptr, 16u, nullptr);
    count_launch();
    return cudaGetLastError();
}

}  // namespace pirb
