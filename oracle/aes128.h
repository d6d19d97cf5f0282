/* oracle/aes128.h — AES-128 for the CPU oracle (TEST INFRASTRUCTURE, never shipped). See aes128.c. */
#ifndef ORC_AES128_H
#define ORC_AES128_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
void orc_aes128_expand_key(const uint8_t key[16], uint8_t rk[176]);
void orc_aes128_encrypt(const uint8_t rk[176], const uint8_t in[16], uint8_t out[16]);
/* out[16i..16i+16) = AES_{k_i}(x) ^ x for i = 0..2 — prf() with numBlocks = 3 (dpf/common.go:62-75) */
void orc_mmo3(const uint8_t rk[3][176], const uint8_t x[16], uint8_t out[48]);
const uint8_t* orc_aes_sbox(void);
int orc_aes_using_aesni(void);
void orc_aes_force_portable(int on);
/* 0 when the portable and the dispatched implementation both match FIPS-197 App. B and C.1 */
int orc_aes_selftest(void);
#ifdef __cplusplus
}
#endif
#endif
