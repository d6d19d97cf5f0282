// keyword_kernel.cu — DPF evaluation at arbitrary points (keyword queries, db.go:132-135: numBits = 32 and
// the point is Keywords[row]; also the batched dpf API) with prefix sharing between consecutive points.
//
// The reference walks all numBits levels for every row.  Rows of a keyword database are (usually) sorted by
// keyword, so consecutive points share a long prefix of their root-to-leaf paths.  Every thread takes a run
// of consecutive points and remembers BOTH children of every node on its current path (thread-private
// local memory, L1/L2 resident); the walk for the next point restarts below the last common node:
// numBits - 1 - cp node expansions instead of numBits, where cp = number of common leading bits (about 4
// instead of 32 for 2^28 sorted keywords in a 32-bit domain).  Results are identical for any order of
// points; only the speed depends on it.  Each lane runs its own (point, level) state machine, one node
// expansion per step, so lanes with short and long restarts do not wait for each other.
#include "kernels.cuh"
#include "aes_unrolled.cuh"

namespace pirb {

extern __shared__ __align__(16) uint8_t g_kwsmem[];

constexpr int kKwThreads = 512;

__global__ void __launch_bounds__(kKwThreads, 1)
keyword_walk_kernel(const __grid_constant__ QueryConst qc, const uint64_t* __restrict__ xs, uint64_t x_first, uint64_t n,
                    uint32_t run_len, int64_t* __restrict__ values, uint8_t* __restrict__ bits,
                    const uint32_t* __restrict__ t0) {
    if (smem_addr(g_kwsmem) > kTableBase) __trap();
    fill_tables(t0);
    __syncthreads();
    const uint32_t lc = lane_const();
    const uint32_t nb = qc.num_bits;
    const uint64_t dom_mask = nb >= 64u ? ~0ull : ((1ull << nb) - 1ull);
    // both children of the path node at every level: [level][0 = left, 1 = right]
    uint4 cs[kMaxLevels][2];
    uint8_t ct[kMaxLevels];  // bit 0 = tL, bit 1 = tR
    const uint64_t n_runs = (n + run_len - 1) / run_len;
    // the run loop and the step loop are warp-uniform (vote on a per-lane `active` flag): the lanes of a warp
    // re-converge every step, so a step always expands one node for every lane that still has work
    for (uint64_t run0 = (uint64_t)blockIdx.x * kKwThreads + (threadIdx.x & ~31u); run0 < n_runs;
         run0 += (uint64_t)gridDim.x * kKwThreads) {
        const uint64_t run = run0 + (threadIdx.x & 31u);
        bool active = run < n_runs;
        uint64_t i = active ? run * run_len : 0;
        const uint64_t i_end = i + run_len < n ? i + run_len : n;
        uint64_t x = active ? ((xs ? xs[i] : x_first + i) & dom_mask) : 0;
        uint32_t s[4] = {qc.s_init[0], qc.s_init[1], qc.s_init[2], qc.s_init[3]};
        uint32_t t = qc.t_init, lvl = 0;
        while (__any_sync(0xffffffffu, active)) {
            if (active && lvl < nb) {
                const uint32_t cw[4] = {qc.cw[lvl][0], qc.cw[lvl][1], qc.cw[lvl][2], qc.cw[lvl][3]};
                const uint32_t cwtl = (uint32_t)(qc.cw_tl >> lvl) & 1u, cwtr = (uint32_t)(qc.cw_tr >> lvl) & 1u;
                uint32_t a[3][4];
                aes3_main_rounds(qc, s, lc, a);
                Children ch;
                inner_tail(qc, a, s, t, cw, cwtl, cwtr, lc, ch);
                cs[lvl][0] = make_uint4(ch.sl[0], ch.sl[1], ch.sl[2], ch.sl[3]);
                cs[lvl][1] = make_uint4(ch.sr[0], ch.sr[1], ch.sr[2], ch.sr[3]);
                ct[lvl] = (uint8_t)(ch.tl | (ch.tr << 1));
                const uint32_t bit = (uint32_t)(x >> (nb - 1u - lvl)) & 1u;  // getBit, dpf/common.go:56-58
#pragma unroll
                for (int c = 0; c < 4; c++) s[c] = bit ? ch.sr[c] : ch.sl[c];
                t = bit ? ch.tr : ch.tl;
                lvl++;
            }
            if (active && lvl == nb) {
                // leaf: emit, then restart the walk for the next point below the last common node
                if (values) values[i] = leaf_value(s[0], s[1], t, qc.final_cw, qc.share_number);
                if (bits) bits[i] = (uint8_t)leaf_select_bit(s[0], s[1], t, qc.fcw_parity);
                i++;
                if (i >= i_end) {
                    active = false;
                } else {
                    const uint64_t xn = (xs ? xs[i] : x_first + i) & dom_mask;
                    const uint64_t diff = x ^ xn;
                    x = xn;
                    if (diff != 0) {  // (diff == 0: the same point again, the leaf in (s, t) is still right)
                        const uint32_t cp = nb - (64u - (uint32_t)__clzll((long long)diff));  // common leading bits
                        const uint32_t bit = (uint32_t)(xn >> (nb - 1u - cp)) & 1u;
                        const uint4 sv = cs[cp][bit];
                        s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
                        t = (ct[cp] >> bit) & 1u;
                        lvl = cp + 1u;
                    }
                }
            }
        }
    }
}

cudaError_t configure_keyword_kernel() {
    return cudaFuncSetAttribute(keyword_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTableEnd);
}

// values and/or bits for n points; xs == nullptr means x = x_first + i.
cudaError_t launch_keyword_walk(const QueryConst& qc, const uint64_t* xs, uint64_t x_first, uint64_t n, int64_t* values,
                                uint8_t* bits, const uint32_t* d_t0, uint32_t smem_base, int num_sms, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    // one run per thread: the points are split evenly over all resident threads, so every SM gets the same
    // number of points and the first (full) walk of a run is amortised over as many points as possible
    const uint64_t threads = (uint64_t)num_sms * kKwThreads;
    uint64_t run_len = (n + threads - 1) / threads;
    if (run_len < 1) run_len = 1;
    const uint64_t n_runs = (n + run_len - 1) / run_len;
    uint64_t blocks = (n_runs + kKwThreads - 1) / kKwThreads;
    if (blocks > (uint64_t)num_sms) blocks = (uint64_t)num_sms;
    keyword_walk_kernel<<<(unsigned)blocks, kKwThreads, kTableEnd - smem_base, st>>>(qc, xs, x_first, n, (uint32_t)run_len, values,
                                                                                     bits, d_t0);
    count_launch();
    return cudaGetLastError();
}

}  // namespace pirb
