/* cabi_smoke.c — plain C99 client of include/pir_b200.h, the way a cgo binding sees the library: no C++, no Python.
 * Runs one committed golden case end to end (tests/golden/dpf_pir_vectors.json, written into golden_case.h by the
 * test that builds this file): client key generation from the 97-byte tape, database upload, both servers' answers
 * through pir_answer, pir_expand_bits + pir_answer_with_bits, the row-sharded pir_server, Recover.
 * Reference path mirrored: db_test.go:18-64 (TestSharedQuery).  Exit code 0 = every byte matches. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "pir_b200.h"
#include "golden_case.h" /* G_DB_SIZE, G_SLOT_BYTES, G_GROUP, G_INDEX, G_NUM_BITS, G_TAPE[], G_DB[], G_ANSWER0[], G_ANSWER1[], G_QUERY0[] */

#define CHECK(call)                                                              \
    do {                                                                         \
        int rc_ = (call);                                                        \
        if (rc_ != PIR_OK) {                                                     \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc_, pir_last_error());     \
            return 1;                                                            \
        }                                                                        \
    } while (0)

int main(void) {
    const size_t ans = (size_t)G_GROUP * G_SLOT_BYTES;
    pir_query q[2];
    uint8_t out[2][4096], via_bits[4096], srv_out[4096];
    uint64_t first = 0, rows = 0;
    pir_db* db = NULL;
    pir_server* srv = NULL;
    int devs[2] = {0, 0};
    if (pir_abi_version() != PIR_B200_ABI_VERSION) return 2;
    if (sizeof(pir_query) != 1256) return 3;
    if (pir_device_count() <= 0) {
        fprintf(stderr, "no CUDA device (the library has no CPU path)\n");
        return 4;
    }
    /* client: ClientInitialize + GenerateTwoServer on the tape (dpf/client.go:17-150, query.go:52-101) */
    CHECK(pir_client_gen_two_server(G_NUM_BITS, G_INDEX, 1, G_TAPE, &q[0], &q[1]));
    q[0].group_size = q[1].group_size = G_GROUP;
    if (memcmp(&q[0], G_QUERY0, sizeof(pir_query)) != 0) {
        fprintf(stderr, "generated key differs from the golden key\n");
        return 5;
    }
    /* the shares travel to their servers in the wire encoding; each server answers the DECODED share */
    for (int s = 0; s < 2; s++) {
        uint8_t wire[PIR_WIRE_HEADER_BYTES + PIR_CW_BYTES * PIR_MAX_NUM_BITS];
        size_t n = 0;
        pir_query back;
        CHECK(pir_query_encode(&q[s], wire, sizeof wire, &n));
        if (n != pir_query_wire_size(G_NUM_BITS)) return 8;
        CHECK(pir_query_decode(wire, n, &back));
        if (memcmp(&back, &q[s], sizeof(pir_query)) != 0) {
            fprintf(stderr, "decoded share differs from the encoded one\n");
            return 8;
        }
        q[s] = back;
    }
    /* server: Database{Slots} -> device handle, then PrivateSecretSharedQuery for both shares (db.go:68-72) */
    CHECK(pir_db_create(0, G_DB_SIZE, G_SLOT_BYTES, 0, G_DB_SIZE, &db));
    CHECK(pir_db_upload(db, 0, G_DB, G_DB_SIZE));
    for (int s = 0; s < 2; s++) CHECK(pir_answer(db, &q[s], out[s], ans));
    if (memcmp(out[0], G_ANSWER0, ans) || memcmp(out[1], G_ANSWER1, ans)) {
        fprintf(stderr, "answer differs from the golden answer\n");
        return 6;
    }
    /* ExpandSharedQuery + ...WithExpandedBits (db.go:75-174) */
    CHECK(pir_db_rows(db, G_GROUP, &first, &rows));
    {
        uint8_t* bits = (uint8_t*)malloc(rows ? rows : 1);
        CHECK(pir_expand_bits(db, &q[0], bits, rows));
        CHECK(pir_answer_with_bits(db, G_GROUP, bits, rows, via_bits, ans));
        free(bits);
    }
    if (memcmp(via_bits, G_ANSWER0, ans)) return 7;
    /* the same database row-sharded over two handles of one server */
    CHECK(pir_server_create(devs, 2, G_DB_SIZE, G_SLOT_BYTES, G_GROUP, &srv));
    CHECK(pir_server_upload(srv, 0, G_DB, G_DB_SIZE));
    CHECK(pir_server_answer(srv, &q[1], srv_out, ans));
    if (memcmp(srv_out, G_ANSWER1, ans)) return 8;
    /* client: Recover (query.go:278-297) */
    pir_xor_combine(out[0], out[1], ans);
    if (memcmp(out[0], G_DB + (size_t)G_INDEX * G_GROUP * G_SLOT_BYTES, ans)) return 9;
    pir_server_destroy(srv);
    pir_db_destroy(db);
    printf("cabi_smoke ok: %s\n", G_NAME);
    return 0;
}
