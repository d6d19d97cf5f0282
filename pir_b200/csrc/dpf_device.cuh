// dpf_device.cuh — device-side building blocks of the DPF tree expansion for sm_100a.
//
// What is computed (reference lines): the PRG of dpf/common.go:62-75 (three fixed-key AES-128 in
// Matyas-Meyer-Oseas mode), one node expansion of (*Dpf).Evaluate2P's loop body
// (dpf/server.go:62-93) for BOTH children at once, and the leaf rule of dpf/server.go:95-100 +
// db.go:142 reduced to the parity it is used for.
//
// How: AES with four T-tables in shared memory, each replicated once per lane (entry x of table
// pair P lives at P*64K + x*256 + half*128 + lane*4), so that the 32 lanes of a warp always hit 32
// different banks: every lookup is one conflict-free LDS.  The tables sit at shared addresses
// [64K, 192K); a lookup address is built with ONE PRMT from the state word and a per-lane constant
// (bytes: lane*4 | lane*4+128 | 0x01 | 0x00), so a table lookup costs PRMT + LDS and nothing else.
// Round keys and correction words are read from the kernel-parameter constant bank.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace pirb {

constexpr uint32_t kTableBase = 65536;        // shared-window address of table pair 0 (T0|T1)
constexpr uint32_t kTableEnd = 3 * 65536;     // end of pair 1 (T2|T3)
constexpr int kMaxLevels = 64;

// Per-query constants, passed by value as a __grid_constant__ kernel parameter.
struct QueryConst {
    uint32_t rk[3][44];          // AES-128 round keys of PrfKeys[0..2], little-endian words
    uint32_t cw[kMaxLevels][4];  // sCW of each level (Key2P.CW[i][0:16])
    uint64_t cw_tl;              // bit i = CW[i][16] & 1
    uint64_t cw_tr;              // bit i = CW[i][17] & 1
    uint32_t s_init[4];
    uint32_t t_init;
    uint32_t fcw_parity;         // FinalCW & 1
    uint32_t num_bits;
    uint32_t share_number;
    int64_t final_cw;
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ uint32_t lds0(uint32_t a) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds1(uint32_t a) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+65536];" : "=r"(v) : "r"(a));
    return v;
}

__device__ __forceinline__ uint32_t lane_const() {
    uint32_t l4 = (threadIdx.x & 31u) * 4u;
    return l4 | ((l4 + 128u) << 8) | (1u << 16);
}

// Fill the lane-replicated tables.  t0 points at the 256-word base table T0 (global/constant).
// All threads of the CTA call this; the caller __syncthreads() afterwards.
__device__ __forceinline__ void fill_tables(const uint32_t* __restrict__ t0) {
    // word index w in [0, 32768): pair = w >> 14, x = (w >> 6) & 255, half = (w >> 5) & 1, lane = w & 31;
    // four consecutive words hold the same value, so every thread stores 16 bytes at a time
    for (uint32_t q = threadIdx.x; q < 8192u; q += blockDim.x) {
        const uint32_t w = q * 4u;
        uint32_t pair = w >> 14, x = (w >> 6) & 255u, half = (w >> 5) & 1u;
        uint32_t v = t0[x];
        uint32_t rot = (pair * 2u + half) * 8u;  // T1 = rotl(T0,8), T2 = rotl(T0,16), T3 = rotl(T0,24)
        v = __funnelshift_l(v, v, rot);
        asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(kTableBase + w * 4u), "r"(v) : "memory");
    }
}

// One middle round for one block: n = MixColumns(ShiftRows(SubBytes(a))) ^ rk
__device__ __forceinline__ void aes_round(const uint32_t a[4], uint32_t n[4], const uint32_t* rk,
                                          uint32_t lc) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
        uint32_t v0 = lds0(__byte_perm(a[c], lc, 0x7604));
        uint32_t v1 = lds0(__byte_perm(a[(c + 1) & 3], lc, 0x7615));
        uint32_t v2 = lds1(__byte_perm(a[(c + 2) & 3], lc, 0x7624));
        uint32_t v3 = lds1(__byte_perm(a[(c + 3) & 3], lc, 0x7635));
        n[c] = v0 ^ v1 ^ v2 ^ v3 ^ rk[c];
    }
}

// Final round output word c (SubBytes + ShiftRows + AddRoundKey), XORed with `mmo` (the PRG input
// word, Matyas-Meyer-Oseas feed-forward).
__device__ __forceinline__ uint32_t aes_last_word(const uint32_t a[4], int c, uint32_t rkw,
                                                  uint32_t mmo, uint32_t lc) {
    uint32_t v0 = lds1(__byte_perm(a[c], lc, 0x7604));            // T2[.]: S in byte 0
    uint32_t v1 = lds1(__byte_perm(a[(c + 1) & 3], lc, 0x7615));  // T3[.]: S in byte 1
    uint32_t v2 = lds0(__byte_perm(a[(c + 2) & 3], lc, 0x7624));  // T0[.]: S in byte 2
    uint32_t v3 = lds0(__byte_perm(a[(c + 3) & 3], lc, 0x7635));  // T1[.]: S in byte 3
    uint32_t lo = __byte_perm(v0, v1, 0x7650);
    uint32_t hi = __byte_perm(v2, v3, 0x7210);
    return __byte_perm(lo, hi, 0x7610) ^ rkw ^ mmo;
}

struct Children {
    uint32_t sl[4], sr[4];
    uint32_t tl, tr;  // 0/1
};

// PRG + correction of one node: both children (dpf/server.go:68-92 for xBit = 0 and 1).
template <int ROUND_UNROLL>
__device__ __forceinline__ void expand_node(const QueryConst& qc, const uint32_t s[4], uint32_t t,
                                            uint32_t level, uint32_t lc, Children& out) {
    uint32_t a[3][4], n[3][4];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int c = 0; c < 4; c++) a[k][c] = s[c] ^ qc.rk[k][c];
#pragma unroll ROUND_UNROLL
    for (int r = 1; r <= 9; r++) {
#pragma unroll
        for (int k = 0; k < 3; k++) aes_round(a[k], n[k], &qc.rk[k][4 * r], lc);
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) a[k][c] = n[k][c];
    }
    uint32_t b0[4], b1[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        b0[c] = aes_last_word(a[0], c, qc.rk[0][40 + c], s[c], lc);
        b1[c] = aes_last_word(a[1], c, qc.rk[1][40 + c], s[c], lc);
    }
    uint32_t b2 = aes_last_word(a[2], 0, qc.rk[2][40], s[0], lc);  // only bytes 0,1 are used
    uint32_t m = 0u - t;
    uint32_t tl = b1[0] & 1u;          // out[16]
    uint32_t tr = (b2 >> 8) & 1u;      // out[33]
#pragma unroll
    for (int c = 0; c < 4; c++) {
        uint32_t nxt = (c < 3) ? b1[c + 1] : b2;
        uint32_t sr = __funnelshift_r(b1[c], nxt, 8);  // out[17+4c .. 21+4c)
        uint32_t cwc = qc.cw[level][c] & m;
        out.sl[c] = b0[c] ^ cwc;
        out.sr[c] = sr ^ cwc;
    }
    out.tl = tl ^ (t & (uint32_t)(qc.cw_tl >> level) & 1u);
    out.tr = tr ^ (t & (uint32_t)(qc.cw_tr >> level) & 1u);
}

// Selection bit of a leaf (s = first 8 seed bytes as two LE words):
//   res = +-(Varint(s[:8]) + t*FinalCW), bit = (res even)            dpf/server.go:95-100, db.go:142
// Varint's parity is bit0^bit1 of s[0] when some byte of s[0..7] is < 0x80, else 0.
__device__ __forceinline__ uint32_t leaf_select_bit(uint32_t w0, uint32_t w1, uint32_t t,
                                                    uint32_t fcw_parity) {
    uint32_t ok = ((w0 & w1 & 0x80808080u) != 0x80808080u) ? 1u : 0u;
    uint32_t par = ok & (w0 ^ (w0 >> 1)) & 1u;
    par ^= t & fcw_parity;
    return par ^ 1u;
}

// Full Varint decode for the dpf API (Evaluate2P returns the integer, not its parity).
__device__ __forceinline__ int64_t leaf_value(uint32_t w0, uint32_t w1, uint32_t t, int64_t final_cw,
                                              uint32_t share_number) {
    uint64_t bytes = ((uint64_t)w1 << 32) | w0;
    uint64_t acc = 0, ux = 0;
    bool done = false;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint64_t b = (bytes >> (8 * j)) & 0xffu;
        if (!done) {
            if (b < 0x80u) {
                ux = acc | (b << (7 * j));
                done = true;
            } else {
                acc |= (b & 0x7fu) << (7 * j);
            }
        }
    }
    if (!done) ux = 0;
    uint64_t x = ux >> 1;
    if (ux & 1u) x = ~x;
    uint64_t v = x + (uint64_t)t * (uint64_t)final_cw;
    if (share_number != 0) v = 0ull - v;
    return (int64_t)v;
}

}  // namespace pirb
