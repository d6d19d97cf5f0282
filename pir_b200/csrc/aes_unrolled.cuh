// aes_unrolled.cuh — the hot-loop form of the PRG node expansion (dpf/common.go:62-75 +
// dpf/server.go:68-92): AES rounds unrolled (kRoundUnroll* below) so that round-key words are constant-bank
// operands of the LOP3 that consumes them, and only the table lookups whose output bytes are actually used
// downstream are issued:
//
//   block 0 (sL):            rounds 1-9 full, last round 16 lookups (8 at a leaf parent)
//   block 1 (tL, sR[0:15]):  rounds 1-9 full, last round 16 lookups (9 at a leaf parent)
//   block 2 (sR[15], tR):    rounds 1-8 full, round 9 columns 0,1 only (column 1 at a leaf parent),
//                            last round 2 lookups (1 at a leaf parent)
//
// 458 lookups for an inner node, 438 for a leaf parent (a full 3-block PRG is 480).
#pragma once
#include "dpf_device.cuh"

namespace pirb {

// How far the middle rounds are unrolled.  Fully unrolled, every round-key word is a constant-bank operand
// (LDCU into a uniform register) but the depth-first loop is ~27 KB of code; the four tree warps of a
// sub-partition sit at different places in it and stall on instruction fetch when it does not fit (the
// even-FinalCW loop, with block 2 under a condition, measured no_instruction 0.20 per issue).  Partly rolled,
// round keys come through per-thread LDC.64 instead.  Measured on 2^24 x 256 B and its shards
// (profiles/r1_round_unroll.txt): even-FinalCW keys are fastest with three iterations of three rounds (1.6 % on
// the whole database, 3-5 % on the shards of 2-8 GPUs); odd-FinalCW keys (lookup pipe 92 % busy) with two
// iterations of four (0.4 % and 2 %) now that every shard runs one task per CTA whose lo warps leave the loop
// early -- with sixteen warps in the loop all the time the full unroll was ahead.
#ifndef PIRB_ROUND_UNROLL_ODD
#define PIRB_ROUND_UNROLL_ODD 4
#endif
#ifndef PIRB_ROUND_UNROLL_EVEN
#define PIRB_ROUND_UNROLL_EVEN 3
#endif
constexpr int kRoundUnrollOdd = PIRB_ROUND_UNROLL_ODD, kRoundUnrollEven = PIRB_ROUND_UNROLL_EVEN;

// state after round 9 for blocks 0,1 and after round 8 for block 2
__device__ __forceinline__ void aes3_main_rounds(const QueryConst& qc, const uint32_t s[4], uint32_t lc,
                                                 uint32_t a[3][4]) {
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int c = 0; c < 4; c++) a[k][c] = s[c] ^ qc.rk[k][c];
#pragma unroll kRoundUnrollOdd
    for (int r = 1; r <= 8; r++) {
        uint32_t n[3][4];
#pragma unroll
        for (int k = 0; k < 3; k++) aes_round(a[k], n[k], &qc.rk[k][4 * r], lc);
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) a[k][c] = n[k][c];
    }
    {
        uint32_t n[2][4];
#pragma unroll
        for (int k = 0; k < 2; k++) aes_round(a[k], n[k], &qc.rk[k][36], lc);
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) a[k][c] = n[k][c];
    }
}

// one column of a middle round
__device__ __forceinline__ uint32_t aes_round_col(const uint32_t a[4], int c, uint32_t rkw, uint32_t lc) {
    uint32_t v0 = lds0(__byte_perm(a[c], lc, 0x7604));
    uint32_t v1 = lds0(__byte_perm(a[(c + 1) & 3], lc, 0x7615));
    uint32_t v2 = lds1(__byte_perm(a[(c + 2) & 3], lc, 0x7624));
    uint32_t v3 = lds1(__byte_perm(a[(c + 3) & 3], lc, 0x7635));
    return v0 ^ v1 ^ v2 ^ v3 ^ rkw;
}

// S-box of byte 0 of x, delivered in byte 0 (T2 = [s,3s,2s,s]); of byte 1 of x in byte 1 (T3 = [s,s,3s,2s])
__device__ __forceinline__ uint32_t sbox_b0(uint32_t x, uint32_t lc) { return lds1(__byte_perm(x, lc, 0x7604)); }
__device__ __forceinline__ uint32_t sbox_b1(uint32_t x, uint32_t lc) { return lds1(__byte_perm(x, lc, 0x7615)); }

// Inner node: both children in full, from the state left by aes3_main_rounds.
__device__ __forceinline__ void inner_tail(const QueryConst& qc, const uint32_t a[3][4], const uint32_t s[4],
                                           uint32_t t, const uint32_t cw[4], uint32_t cwtl, uint32_t cwtr,
                                           uint32_t lc, Children& out) {
    const uint32_t n20 = aes_round_col(a[2], 0, qc.rk[2][36], lc);
    const uint32_t n21 = aes_round_col(a[2], 1, qc.rk[2][37], lc);
    uint32_t b0[4], b1[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        b0[c] = aes_last_word(a[0], c, qc.rk[0][40 + c], s[c], lc);
        b1[c] = aes_last_word(a[1], c, qc.rk[1][40 + c], s[c], lc);
    }
    // block 2, bytes 0 and 1 of word 0 only
    const uint32_t u0 = sbox_b0(n20, lc), u1 = sbox_b1(n21, lc);
    const uint32_t b2 = ((u0 & 0xffu) | (u1 & 0xff00u)) ^ qc.rk[2][40] ^ s[0];
    const uint32_t m = 0u - t;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint32_t nxt = (c < 3) ? b1[c + 1] : b2;
        const uint32_t cwc = cw[c] & m;
        out.sl[c] = b0[c] ^ cwc;
        out.sr[c] = __funnelshift_r(b1[c], nxt, 8) ^ cwc;
    }
    out.tl = (b1[0] & 1u) ^ (t & cwtl);
    out.tr = ((b2 >> 8) & 1u) ^ (t & cwtr);
}

// Leaf parent: only what the two selection bits need (8 seed bytes and t of each child).
// Returns bit0 = selection bit of the left leaf, bit1 = of the right leaf.
__device__ __forceinline__ uint32_t leaf_parent_tail(const QueryConst& qc, const uint32_t a[3][4],
                                                     const uint32_t s[4], uint32_t t, const uint32_t cw[4],
                                                     uint32_t cwtl, uint32_t cwtr, uint32_t lc) {
    const uint32_t n21 = aes_round_col(a[2], 1, qc.rk[2][37], lc);
    const uint32_t l0 = aes_last_word(a[0], 0, qc.rk[0][40], s[0], lc);
    const uint32_t l1 = aes_last_word(a[0], 1, qc.rk[0][41], s[1], lc);
    const uint32_t r0 = aes_last_word(a[1], 0, qc.rk[1][40], s[0], lc);
    const uint32_t r1 = aes_last_word(a[1], 1, qc.rk[1][41], s[1], lc);
    const uint32_t r2 = sbox_b0(a[1][2], lc) ^ qc.rk[1][42] ^ s[2];          // byte 0 valid
    const uint32_t b2 = sbox_b1(n21, lc) ^ qc.rk[2][40] ^ s[0];              // byte 1 valid
    const uint32_t m = 0u - t;
    const uint32_t c0 = cw[0] & m, c1 = cw[1] & m;
    const uint32_t sl0 = l0 ^ c0, sl1 = l1 ^ c1;
    const uint32_t sr0 = __funnelshift_r(r0, r1, 8) ^ c0;
    const uint32_t sr1 = __funnelshift_r(r1, r2, 8) ^ c1;
    const uint32_t tl = (r0 & 1u) ^ (t & cwtl);
    const uint32_t tr = ((b2 >> 8) & 1u) ^ (t & cwtr);
    return leaf_select_bit(sl0, sl1, tl, qc.fcw_parity) | (leaf_select_bit(sr0, sr1, tr, qc.fcw_parity) << 1);
}

// ---- leaf parents of a key whose FinalCW is even ------------------------------------------------------
// The selection bit is 1 ^ parity(Varint(s[:8])) ^ (t & FinalCW & 1)  (dpf/server.go:95-100, db.go:142).
// When FinalCW is even the leaf's control bit t does not enter it, so at a leaf parent neither tL nor tR
// is needed: block 2 of the PRG (whose only outputs are sR[15] and tR) is skipped altogether and blocks 0
// and 1 deliver the 8 seed bytes of each child.  305 lookups instead of 438.  (Half of all keys: the
// parity of FinalCW = b - Varint(s0) + Varint(s1) is a coin flip, dpf/client.go:127-141.)
// The same instruction stream as aes3_main_rounds with block 2 under a (warp-uniform) condition: inner nodes
// and even-FinalCW leaf parents then share one copy of the unrolled rounds.  With a separate two-block copy
// the depth-first loop was ~42 KB of code and the warps of a sub-partition, all at different places in it,
// stalled on instruction fetch (ncu: no_instruction 0.53 per issued instruction against 0.05 for the
// ~27 KB loop of odd-FinalCW keys).
__device__ __forceinline__ void aes3_main_rounds_cond(const QueryConst& qc, const uint32_t s[4], uint32_t lc,
                                                      uint32_t a[3][4], bool with_block2) {
#pragma unroll
    for (int k = 0; k < 2; k++)
#pragma unroll
        for (int c = 0; c < 4; c++) a[k][c] = s[c] ^ qc.rk[k][c];
    if (with_block2) {
#pragma unroll
        for (int c = 0; c < 4; c++) a[2][c] = s[c] ^ qc.rk[2][c];
    }
#pragma unroll kRoundUnrollEven
    for (int r = 1; r <= 9; r++) {
        uint32_t n[3][4];
#pragma unroll
        for (int k = 0; k < 2; k++) aes_round(a[k], n[k], &qc.rk[k][4 * r], lc);
        if (r <= 8 && with_block2) {
            aes_round(a[2], n[2], &qc.rk[2][4 * r], lc);
#pragma unroll
            for (int c = 0; c < 4; c++) a[2][c] = n[2][c];
        }
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) a[k][c] = n[k][c];
    }
}

// Returns bit0 = selection bit of the left leaf, bit1 = of the right leaf (FinalCW even).
__device__ __forceinline__ uint32_t leaf_parent_tail_even(const QueryConst& qc, const uint32_t a[3][4],
                                                          const uint32_t s[4], uint32_t t, const uint32_t cw[4],
                                                          uint32_t lc) {
    const uint32_t l0 = aes_last_word(a[0], 0, qc.rk[0][40], s[0], lc);
    const uint32_t l1 = aes_last_word(a[0], 1, qc.rk[0][41], s[1], lc);
    const uint32_t r0 = aes_last_word(a[1], 0, qc.rk[1][40], s[0], lc);
    const uint32_t r1 = aes_last_word(a[1], 1, qc.rk[1][41], s[1], lc);
    const uint32_t r2 = sbox_b0(a[1][2], lc) ^ qc.rk[1][42] ^ s[2];          // byte 0 valid
    const uint32_t m = 0u - t;
    const uint32_t c0 = cw[0] & m, c1 = cw[1] & m;
    const uint32_t sr0 = __funnelshift_r(r0, r1, 8) ^ c0;
    const uint32_t sr1 = __funnelshift_r(r1, r2, 8) ^ c1;
    return leaf_select_bit(l0 ^ c0, l1 ^ c1, 0u, 0u) | (leaf_select_bit(sr0, sr1, 0u, 0u) << 1);
}

// ---- one child of a leaf parent (keyword_kernel.cu) --------------------------------------------------------
// A keyword walk needs exactly ONE child of its deepest node, and of that child only the 8 seed bytes the selection bit is
// computed from plus -- for keys with an odd FinalCW, or when the full Evaluate2P value is wanted -- its control bit:
//   left  child: sL[0:8] = block 0 words 0,1;  tL = block 1 byte 0            152 (+145) lookups
//   right child: sR[0:8] = block 1 bytes 1..8; tR = block 2 byte 1            153 (+133) lookups
// against 458 for both children in full.  All lanes of the calling warp must want the same side (the caller batches).
template <int NR>
__device__ __forceinline__ void aes_block_rounds(const QueryConst& qc, int k, const uint32_t s[4], uint32_t lc, uint32_t a[4]) {
#pragma unroll
    for (int c = 0; c < 4; c++) a[c] = s[c] ^ qc.rk[k][c];
#pragma unroll 3
    for (int r = 1; r <= NR; r++) {
        uint32_t n[4];
        aes_round(a, n, &qc.rk[k][4 * r], lc);
#pragma unroll
        for (int c = 0; c < 4; c++) a[c] = n[c];
    }
}
// two blocks side by side (two independent dependency chains): rounds 1..8 of blocks k0 and k1, then round 9 of k0 only if R9_BOTH is false
template <bool R9_BOTH>
__device__ __forceinline__ void aes_two_blocks(const QueryConst& qc, int k0, int k1, const uint32_t s[4], uint32_t lc, uint32_t a[4],
                                               uint32_t b[4]) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
        a[c] = s[c] ^ qc.rk[k0][c];
        b[c] = s[c] ^ qc.rk[k1][c];
    }
#pragma unroll 2
    for (int r = 1; r <= 8; r++) {
        uint32_t n[4], m[4];
        aes_round(a, n, &qc.rk[k0][4 * r], lc);
        aes_round(b, m, &qc.rk[k1][4 * r], lc);
#pragma unroll
        for (int c = 0; c < 4; c++) {
            a[c] = n[c];
            b[c] = m[c];
        }
    }
    {
        uint32_t n[4];
        aes_round(a, n, &qc.rk[k0][36], lc);
#pragma unroll
        for (int c = 0; c < 4; c++) a[c] = n[c];
    }
    if (R9_BOTH) {
        uint32_t m[4];
        aes_round(b, m, &qc.rk[k1][36], lc);
#pragma unroll
        for (int c = 0; c < 4; c++) b[c] = m[c];
    }
}

// (w0, w1) = the child's first 8 seed bytes, tc = its control bit (0 when need_t is false: the caller must not use it then)
__device__ __forceinline__ void leaf_child_left(const QueryConst& qc, const uint32_t s[4], uint32_t t, uint32_t cw0, uint32_t cw1,
                                                uint32_t cwtl, bool need_t, uint32_t lc, uint32_t& w0, uint32_t& w1, uint32_t& tc) {
    const uint32_t m = 0u - t;
    uint32_t a[4], b[4];
    tc = 0u;
    if (need_t) {
        aes_two_blocks<true>(qc, 0, 1, s, lc, a, b);
        const uint32_t r0 = sbox_b0(b[0], lc) ^ qc.rk[1][40] ^ s[0];  // byte 0 of block 1's output word 0
        tc = (r0 & 1u) ^ (t & cwtl);
    } else {
        aes_block_rounds<9>(qc, 0, s, lc, a);
    }
    w0 = aes_last_word(a, 0, qc.rk[0][40], s[0], lc) ^ (cw0 & m);
    w1 = aes_last_word(a, 1, qc.rk[0][41], s[1], lc) ^ (cw1 & m);
}
__device__ __forceinline__ void leaf_child_right(const QueryConst& qc, const uint32_t s[4], uint32_t t, uint32_t cw0, uint32_t cw1,
                                                 uint32_t cwtr, bool need_t, uint32_t lc, uint32_t& w0, uint32_t& w1, uint32_t& tc) {
    const uint32_t m = 0u - t;
    uint32_t a[4], b[4];
    tc = 0u;
    if (need_t) {
        aes_two_blocks<false>(qc, 1, 2, s, lc, a, b);  // block 1 after round 9, block 2 after round 8
        const uint32_t n21 = aes_round_col(b, 1, qc.rk[2][37], lc);
        const uint32_t b2 = sbox_b1(n21, lc) ^ qc.rk[2][40] ^ s[0];  // byte 1 of block 2's output word 0
        tc = ((b2 >> 8) & 1u) ^ (t & cwtr);
    } else {
        aes_block_rounds<9>(qc, 1, s, lc, a);
    }
    const uint32_t r0 = aes_last_word(a, 0, qc.rk[1][40], s[0], lc);
    const uint32_t r1 = aes_last_word(a, 1, qc.rk[1][41], s[1], lc);
    const uint32_t r2 = sbox_b0(a[2], lc) ^ qc.rk[1][42] ^ s[2];      // byte 0 valid
    w0 = __funnelshift_r(r0, r1, 8) ^ (cw0 & m);
    w1 = __funnelshift_r(r1, r2, 8) ^ (cw1 & m);
}

}  // namespace pirb
