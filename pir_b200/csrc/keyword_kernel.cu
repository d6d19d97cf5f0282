// keyword_kernel.cu — DPF evaluation at arbitrary points (keyword queries, db.go:132-135: numBits = 32 and
// the point is Keywords[row]; also the batched dpf API) with prefix sharing between consecutive points, and, for the
// answer path, the row XOR of db.go:91-104 fused into the same kernel.
//
// The reference walks all numBits levels for every row.  Rows of a keyword database are (usually) sorted by
// keyword, so consecutive points share a long prefix of their root-to-leaf paths.  Every thread takes a run of
// consecutive points and walks them like a depth-first traversal of the sparse tree they span: whenever it turns
// LEFT it keeps the right child of that node (a sibling stack: the deepest kKwSmemLevels levels in shared memory,
// one conflict-free column per thread, the rare shallower ones in thread-private memory); the walk for the next
// point restarts from the stacked sibling at the first level where the two points differ: numBits - 1 - cp node
// expansions instead of numBits, where cp = number of common leading bits (about 4.5 instead of 32 for 2^28 sorted
// keywords in a 32-bit domain).  A point that does not continue the traversal (smaller than its predecessor, or
// the sibling it needs was never stacked) restarts at the root: results are identical for ANY order of points, only
// the speed depends on it.  Each lane runs its own (point, level) state machine, one node expansion per step, so
// lanes with short and long restarts do not wait for each other.
//
// FUSED mode (the keyword answer path): when a lane reaches a leaf whose selection bit is set, the warp fetches that
// row cooperatively (lane = 16-byte chunk, 32/c rows per load for rows of c chunks) and XORs it into per-lane
// accumulators: no byte-per-row bit vector goes through HBM and no second kernel runs.
#include "kernels.cuh"
#include "aes_unrolled.cuh"

namespace pirb {

extern __shared__ __align__(16) uint8_t g_kwsmem[];

constexpr int kKwThreads = 512;
constexpr uint32_t kKwLevelBytes = kKwThreads * 16;              // one stacked sibling per thread per level
constexpr uint32_t kKwLowLevels = 7;                             // levels below the tables: [base, 64K)
constexpr uint32_t kKwHighLevels = 4;                            // levels above them: [192K, 224K)
constexpr uint32_t kKwSmemLevels = kKwLowLevels + kKwHighLevels;
constexpr uint32_t kKwAbsHigh = kTableEnd;                                      // absolute shared addresses
constexpr uint32_t kKwAbsCw = kKwAbsHigh + kKwHighLevels * kKwLevelBytes;      // [64][4] u32 correction words
constexpr uint32_t kKwAbsRed = kKwAbsCw + kMaxLevels * 16;                      // [128] u32: the CTA's partial answer
constexpr uint32_t kKwSmemEnd = kKwAbsRed + 512;
static_assert(kKwSmemEnd <= 227u * 1024u, "keyword kernel shared memory");

// shared-memory word address of word w of thread tid's sibling at stack level sl (conflict-free: tid is the fast index)
__device__ __forceinline__ uint32_t kw_sib_addr(uint32_t dyn_base, uint32_t sl, uint32_t w, uint32_t tid) {
    const uint32_t lvl_base = sl < kKwLowLevels ? dyn_base + sl * kKwLevelBytes : kKwAbsHigh + (sl - kKwLowLevels) * kKwLevelBytes;
    return lvl_base + (w * kKwThreads + tid) * 4u;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
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
        if (dyn_base + kKwLowLevels * kKwLevelBytes > kTableBase || dyn_base + dyn < kKwSmemEnd) __trap();
    }
    fill_tables(t0);
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t nb = qc.num_bits;
    // correction words in shared memory: the lanes of a warp sit at different levels, and an indexed constant-bank
    // load serialises over distinct addresses
    for (uint32_t w = tid; w < nb * 4u; w += kKwThreads) sts32(kKwAbsCw + w * 4u, qc.cw[w >> 2][w & 3u]);
    if (FUSED)
        for (uint32_t w = tid; w < 128u; w += kKwThreads) sts32(kKwAbsRed + w * 4u, 0u);
    __syncthreads();
    const uint32_t lc = lane_const();
    const uint64_t dom_mask = nb >= 64u ? ~0ull : ((1ull << nb) - 1ull);
    const uint32_t smem_l0 = nb > kKwSmemLevels ? nb - kKwSmemLevels : 0u;  // first level whose siblings live in shared memory
    uint4 deep[kMaxLevels - kKwSmemLevels];  // siblings of the levels above smem_l0 (thread-private, rarely touched)
    // FUSED: lane = (row slot, 16-byte chunk): rows of c chunks, rpl rows per warp load
    const uint32_t c = FUSED ? row_bytes >> 4 : 1u, rpl = FUSED ? 32u / c : 1u;
    const uint32_t my_slot = lane / c, my_chunk = lane - my_slot * c;
    uint4 acc = make_uint4(0, 0, 0, 0);
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
            if (active && lvl < nb) {
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
            const bool leaf = active && lvl == nb;
            uint32_t sel = 0;
            if (leaf) {
                if (values) values[i] = leaf_value(s[0], s[1], t, qc.final_cw, qc.share_number);
                sel = leaf_select_bit(s[0], s[1], t, qc.fcw_parity);
                if (bits) bits[i] = (uint8_t)sel;
            }
            if (FUSED) {
                // rows selected in this step, fetched by the whole warp: slot k of an iteration takes the k-th of them
                const uint32_t bal = __ballot_sync(0xffffffffu, leaf && sel);
                const uint32_t cnt = __popc(bal);
                for (uint32_t k0 = 0; k0 < cnt; k0 += rpl) {
                    const uint32_t k = k0 + my_slot;
                    const uint32_t src = k < cnt ? __fns(bal, 0u, (int)k + 1) : 0u;
                    const uint32_t ilo = __shfl_sync(0xffffffffu, (uint32_t)i, src & 31u);
                    const uint32_t ihi = __shfl_sync(0xffffffffu, (uint32_t)(i >> 32), src & 31u);
                    if (k < cnt && my_slot < rpl) {
                        const uint64_t row = ((uint64_t)ihi << 32) | ilo;
                        const uint4 v = __ldg(reinterpret_cast<const uint4*>(db + row * row_bytes) + my_chunk);
                        acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
                    }
                }
            }
            if (leaf) {
                // restart the walk for the next point from the sibling stacked where the two points part
                i++;
                if (i >= i_end) {
                    active = false;
                } else {
                    const uint64_t xn = (xs ? xs[i] : x_first + i) & dom_mask;
                    const uint64_t diff = x ^ xn;
                    x = xn;
                    if (diff != 0) {  // (diff == 0: the same point again, the leaf in (s, t) is still right)
                        const uint32_t cp = nb - (64u - (uint32_t)__clzll((long long)diff));  // common leading bits
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
        }
    }
    if (FUSED) {
        if (my_slot < rpl && (acc.x | acc.y | acc.z | acc.w)) {
            const uint32_t ra = kKwAbsRed + my_chunk * 16u;
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
