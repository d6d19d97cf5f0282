/*
 * oracle/pir_oracle.h — CPU restatement of the reference's DPF-PIR answer path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or the timed CPU baseline.
 * Nothing under pir_b200/ links, imports or calls it.
 *
 * Parity status: the reference holds NO golden vectors or known-answer tests for this path (all its
 * tests draw keys from crypto/rand and check round-trip properties: db_test.go:18-64,
 * dpf/dpf_test.go:11-107) and no Go toolchain exists in the build container, so share-level bytes
 * are "parity unpinned" against a Go binary.  They are pinned instead by (1) FIPS-197 AES vectors,
 * (2) the reference's own properties re-run on this oracle (tests/test_oracle_properties.py), and
 * (3) an independent pure-Python restatement that uses the `cryptography` package's AES
 * (oracle/py_restatement.py) agreeing byte-for-byte on the committed vectors in tests/golden/.
 */
#ifndef PIR_ORACLE_H
#define PIR_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#include "../include/pir_b200.h" /* the pir_query POD only */
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_TAPE_BYTES 97 /* 4*16 PRF keys + 17 (SInit0||TInit0 source) + 16 SInit1 */

typedef struct orc_dpf {
    uint8_t rk[3][176];
    uint32_t num_bits;
} orc_dpf;

/* binary.Varint(buf[:8]) (dpf/server.go:95, dpf/client.go:141-142) */
int64_t orc_varint8(const uint8_t s[8]);
/* dpf.ClientInitialize + GenerateTwoServer(a, b) (dpf/client.go:17-150) with crypto/rand replaced by
 * the caller's 97-byte tape, consumed in the reference's order of rand.Read calls. */
void orc_gen_two_server(const uint8_t tape[ORC_TAPE_BYTES], uint32_t num_bits, uint64_t a,
                        uint64_t b, pir_query* q0, pir_query* q1);
/* dpf.ServerInitialize (dpf/server.go:16-46) */
void orc_server_initialize(const pir_query* q, uint32_t num_bits, orc_dpf* f);
/* (*Dpf).Evaluate2P (dpf/server.go:55-101) */
int64_t orc_evaluate2p(const orc_dpf* f, const pir_query* q, uint64_t x);
/* db.go:114-121 */
uint32_t orc_num_bits_for(uint64_t db_size, uint32_t group_size, int is_keyword);
/* (*Database).ExpandSharedQuery (db.go:110-174) for rows [row_begin, row_begin+n_rows): one
 * root-to-leaf Evaluate2P per row, exactly like the reference.  keywords (may be NULL) is indexed by
 * the GLOBAL row number.  nthreads >= 1 splits the rows statically. */
void orc_expand_shared_query(const pir_query* q, uint64_t db_size, const uint64_t* keywords,
                             uint64_t row_begin, uint64_t n_rows, uint8_t* bits, int nthreads);
/* Same bits by GGM tree expansion (index queries only) — the algorithm the GPU uses, on the CPU. */
void orc_expand_tree(const pir_query* q, uint64_t db_size, uint64_t row_begin, uint64_t n_rows,
                     uint8_t* bits, int nthreads);
/* (*Database).PrivateSecretSharedQueryWithExpandedBits (db.go:75-107) + XorSlots (slot.go:19-30).
 * slots = dense slots of rows [row_begin, ...): slot (row_begin+r)*W + c is at
 * slots + (r*W + c)*slot_bytes.  out: W*slot_bytes bytes (zeroed here). */
void orc_answer_with_bits(const uint8_t* slots, uint32_t slot_bytes, uint32_t group_size,
                          const uint8_t* bits, uint64_t n_rows, uint8_t* out);
/* PrivateSecretSharedQuery (db.go:68-72) over a whole dense database. */
void orc_answer(const pir_query* q, const uint8_t* slots, uint64_t db_size, uint32_t slot_bytes,
                const uint64_t* keywords, uint8_t* out, int nthreads, int use_tree);
/* the synthetic database generator shared with the GPU fill kernel (DESIGN.md "pir_synth") */
void orc_synth_fill(uint64_t seed, uint64_t slot_begin, uint64_t n_slots, uint32_t slot_bytes,
                    uint8_t* out);
uint64_t orc_synth_word(uint64_t seed, uint64_t slot, uint32_t w);
/* db.go:75-107 for nq bit vectors ([nq][n_rows] bytes) over rows [row_begin, +n_rows) of the synthetic database,
 * generated on the fly (no 4-8 GiB host copy); out = [nq][group_size*slot_bytes]. */
void orc_answers_over_synth(uint64_t seed, uint32_t slot_bytes, uint32_t group_size, uint64_t row_begin,
                            uint64_t n_rows, const uint8_t* bits, uint32_t nq, uint8_t* out, int nthreads);
/* deterministic byte tape for key generation in tests (splitmix64 stream) */
void orc_tape(uint64_t seed, uint8_t* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif
