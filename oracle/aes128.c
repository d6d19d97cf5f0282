/*
 * oracle/aes128.c — AES-128 encryption for the CPU oracle (TEST INFRASTRUCTURE, never shipped).
 *
 * The reference uses Go's crypto/aes (aes.NewCipher on 16-byte keys, dpf/server.go:21-33,
 * dpf/client.go:24-35; cipher.Block.Encrypt at dpf/common.go:70).  crypto/aes is absent from
 * /root/reference (Go standard library, version unpinned: the tree has no go.mod); it implements
 * FIPS-197, so any conforming AES-128 is bit-identical.  Two implementations live here:
 *   - a portable one built from the FIPS-197 definition (S-box derived from the GF(2^8) inverse and
 *     the affine map at start-up, no literal tables), and
 *   - an AES-NI one used when the host CPU has it (Go uses AES-NI on amd64 too).
 * orc_aes_selftest() checks both against the FIPS-197 Appendix B / C.1 vectors.
 */
#include "aes128.h"
#include <string.h>

static uint8_t g_sbox[256];
static int g_sbox_ready = 0;

static uint8_t gf_mul(uint8_t a, uint8_t b) {
    uint8_t p = 0;
    for (int i = 0; i < 8; i++) {
        if (b & 1) p ^= a;
        uint8_t hi = a & 0x80;
        a = (uint8_t)(a << 1);
        if (hi) a ^= 0x1b; /* x^8 + x^4 + x^3 + x + 1 */
        b >>= 1;
    }
    return p;
}

static void sbox_init(void) {
    if (g_sbox_ready) return;
    for (int x = 0; x < 256; x++) {
        uint8_t inv = 0;
        if (x) {
            /* x^254 = x^-1 in GF(2^8) */
            uint8_t r = 1, base = (uint8_t)x;
            int e = 254;
            while (e) {
                if (e & 1) r = gf_mul(r, base);
                base = gf_mul(base, base);
                e >>= 1;
            }
            inv = r;
        }
        uint8_t s = inv;
        uint8_t rot = inv;
        for (int k = 0; k < 4; k++) {
            rot = (uint8_t)((rot << 1) | (rot >> 7));
            s ^= rot;
        }
        g_sbox[x] = (uint8_t)(s ^ 0x63);
    }
    g_sbox_ready = 1;
}

const uint8_t* orc_aes_sbox(void) {
    sbox_init();
    return g_sbox;
}

void orc_aes128_expand_key(const uint8_t key[16], uint8_t rk[176]) {
    sbox_init();
    memcpy(rk, key, 16);
    uint8_t rcon = 1;
    for (int i = 16; i < 176; i += 4) {
        uint8_t t[4] = {rk[i - 4], rk[i - 3], rk[i - 2], rk[i - 1]};
        if (i % 16 == 0) {
            uint8_t t0 = t[0];
            t[0] = (uint8_t)(g_sbox[t[1]] ^ rcon);
            t[1] = g_sbox[t[2]];
            t[2] = g_sbox[t[3]];
            t[3] = g_sbox[t0];
            rcon = gf_mul(rcon, 2);
        }
        for (int j = 0; j < 4; j++) rk[i + j] = (uint8_t)(rk[i - 16 + j] ^ t[j]);
    }
}

static void encrypt_portable(const uint8_t rk[176], const uint8_t in[16], uint8_t out[16]) {
    uint8_t s[16], t[16];
    for (int i = 0; i < 16; i++) s[i] = (uint8_t)(in[i] ^ rk[i]);
    for (int round = 1; round <= 10; round++) {
        /* SubBytes + ShiftRows: state is column-major, byte i = row i%4, column i/4 */
        for (int c = 0; c < 4; c++)
            for (int r = 0; r < 4; r++) t[4 * c + r] = g_sbox[s[4 * ((c + r) & 3) + r]];
        if (round < 10) {
            for (int c = 0; c < 4; c++) {
                uint8_t a0 = t[4 * c], a1 = t[4 * c + 1], a2 = t[4 * c + 2], a3 = t[4 * c + 3];
                s[4 * c + 0] = (uint8_t)(gf_mul(a0, 2) ^ gf_mul(a1, 3) ^ a2 ^ a3);
                s[4 * c + 1] = (uint8_t)(a0 ^ gf_mul(a1, 2) ^ gf_mul(a2, 3) ^ a3);
                s[4 * c + 2] = (uint8_t)(a0 ^ a1 ^ gf_mul(a2, 2) ^ gf_mul(a3, 3));
                s[4 * c + 3] = (uint8_t)(gf_mul(a0, 3) ^ a1 ^ a2 ^ gf_mul(a3, 2));
            }
        } else {
            memcpy(s, t, 16);
        }
        for (int i = 0; i < 16; i++) s[i] ^= rk[16 * round + i];
    }
    memcpy(out, s, 16);
}

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
#define ORC_HAVE_X86 1
__attribute__((target("aes,sse2"))) static void encrypt_aesni(const uint8_t rk[176],
                                                              const uint8_t in[16],
                                                              uint8_t out[16]) {
    const __m128i* k = (const __m128i*)rk;
    __m128i b = _mm_xor_si128(_mm_loadu_si128((const __m128i*)in), _mm_loadu_si128(k));
    for (int r = 1; r < 10; r++) b = _mm_aesenc_si128(b, _mm_loadu_si128(k + r));
    b = _mm_aesenclast_si128(b, _mm_loadu_si128(k + 10));
    _mm_storeu_si128((__m128i*)out, b);
}
/* three fixed keys applied to one block, Matyas-Meyer-Oseas: out_i = AES_{k_i}(x) ^ x
 * (dpf/common.go:62-75 with numBlocks = 3) */
__attribute__((target("aes,sse2"))) static void mmo3_aesni(const uint8_t rk[3][176],
                                                           const uint8_t x[16], uint8_t out[48]) {
    __m128i in = _mm_loadu_si128((const __m128i*)x);
    const __m128i* k0 = (const __m128i*)rk[0];
    const __m128i* k1 = (const __m128i*)rk[1];
    const __m128i* k2 = (const __m128i*)rk[2];
    __m128i a = _mm_xor_si128(in, _mm_loadu_si128(k0));
    __m128i b = _mm_xor_si128(in, _mm_loadu_si128(k1));
    __m128i c = _mm_xor_si128(in, _mm_loadu_si128(k2));
    for (int r = 1; r < 10; r++) {
        a = _mm_aesenc_si128(a, _mm_loadu_si128(k0 + r));
        b = _mm_aesenc_si128(b, _mm_loadu_si128(k1 + r));
        c = _mm_aesenc_si128(c, _mm_loadu_si128(k2 + r));
    }
    a = _mm_aesenclast_si128(a, _mm_loadu_si128(k0 + 10));
    b = _mm_aesenclast_si128(b, _mm_loadu_si128(k1 + 10));
    c = _mm_aesenclast_si128(c, _mm_loadu_si128(k2 + 10));
    _mm_storeu_si128((__m128i*)(out + 0), _mm_xor_si128(a, in));
    _mm_storeu_si128((__m128i*)(out + 16), _mm_xor_si128(b, in));
    _mm_storeu_si128((__m128i*)(out + 32), _mm_xor_si128(c, in));
}
#endif

static int g_use_aesni = -1;
static int g_force_portable = 0;

void orc_aes_force_portable(int on) {
    g_force_portable = on;
    g_use_aesni = -1;
}

int orc_aes_using_aesni(void) {
    if (g_use_aesni < 0) {
#ifdef ORC_HAVE_X86
        __builtin_cpu_init();
        g_use_aesni = (!g_force_portable && __builtin_cpu_supports("aes")) ? 1 : 0;
#else
        g_use_aesni = 0;
#endif
    }
    return g_use_aesni;
}

void orc_aes128_encrypt(const uint8_t rk[176], const uint8_t in[16], uint8_t out[16]) {
#ifdef ORC_HAVE_X86
    if (orc_aes_using_aesni()) {
        encrypt_aesni(rk, in, out);
        return;
    }
#endif
    encrypt_portable(rk, in, out);
}

void orc_mmo3(const uint8_t rk[3][176], const uint8_t x[16], uint8_t out[48]) {
#ifdef ORC_HAVE_X86
    if (orc_aes_using_aesni()) {
        mmo3_aesni(rk, x, out);
        return;
    }
#endif
    for (int i = 0; i < 3; i++) {
        uint8_t tmp[16];
        encrypt_portable(rk[i], x, tmp);
        for (int j = 0; j < 16; j++) out[16 * i + j] = (uint8_t)(tmp[j] ^ x[j]);
    }
}

/* FIPS-197 Appendix B (key 2b7e..., pt 3243...) and Appendix C.1 (key 00..0f, pt 00..ff). */
int orc_aes_selftest(void) {
    static const uint8_t keyB[16] = {0x2b, 0x7e, 0x15, 0x16, 0x28, 0xae, 0xd2, 0xa6,
                                     0xab, 0xf7, 0x15, 0x88, 0x09, 0xcf, 0x4f, 0x3c};
    static const uint8_t ptB[16] = {0x32, 0x43, 0xf6, 0xa8, 0x88, 0x5a, 0x30, 0x8d,
                                    0x31, 0x31, 0x98, 0xa2, 0xe0, 0x37, 0x07, 0x34};
    static const uint8_t ctB[16] = {0x39, 0x25, 0x84, 0x1d, 0x02, 0xdc, 0x09, 0xfb,
                                    0xdc, 0x11, 0x85, 0x97, 0x19, 0x6a, 0x0b, 0x32};
    static const uint8_t ctC[16] = {0x69, 0xc4, 0xe0, 0xd8, 0x6a, 0x7b, 0x04, 0x30,
                                    0xd8, 0xcd, 0xb7, 0x80, 0x70, 0xb4, 0xc5, 0x5a};
    uint8_t keyC[16], ptC[16], rk[176], out[16];
    for (int i = 0; i < 16; i++) {
        keyC[i] = (uint8_t)i;
        ptC[i] = (uint8_t)(i * 0x11);
    }
    int bad = 0;
    orc_aes128_expand_key(keyB, rk);
    encrypt_portable(rk, ptB, out);
    bad |= memcmp(out, ctB, 16) != 0;
    orc_aes128_encrypt(rk, ptB, out);
    bad |= (memcmp(out, ctB, 16) != 0) << 1;
    orc_aes128_expand_key(keyC, rk);
    encrypt_portable(rk, ptC, out);
    bad |= (memcmp(out, ctC, 16) != 0) << 2;
    orc_aes128_encrypt(rk, ptC, out);
    bad |= (memcmp(out, ctC, 16) != 0) << 3;
    return bad;
}
