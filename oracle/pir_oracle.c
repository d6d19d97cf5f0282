/*
 * oracle/pir_oracle.c — CPU restatement of the reference's two-server DPF-PIR answer path.
 * TEST INFRASTRUCTURE (see pir_oracle.h for who may call it and for the parity status:
 * "parity unpinned" against a Go binary; pinned by FIPS-197 + reference properties + an
 * independent Python restatement).
 *
 * Each function cites the reference lines it follows (paths relative to the reference tree).
 */
#include "pir_oracle.h"
#include "aes128.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- encoding/binary.Varint on an 8-byte slice (dpf/server.go:95) ------------------------------
 * Go's Uvarint walks the bytes; the first byte < 0x80 terminates; if the slice ends first it returns
 * (0, 0).  Varint then zig-zag decodes: x = ux>>1, complemented when ux&1. */
int64_t orc_varint8(const uint8_t s[8]) {
    uint64_t ux = 0, acc = 0;
    unsigned shift = 0;
    for (int j = 0; j < 8; j++) {
        uint8_t b = s[j];
        if (b < 0x80) {
            ux = acc | ((uint64_t)b << shift);
            uint64_t x = ux >> 1;
            if (ux & 1) x = ~x;
            return (int64_t)x;
        }
        acc |= (uint64_t)(b & 0x7f) << shift;
        shift += 7;
    }
    return 0; /* buffer too small: Uvarint returns 0 */
}

/* getBit(x, N-numBits+i+1, N) == bit (numBits-1-i) of x (dpf/common.go:56-58, dpf/server.go:63-66) */
static inline uint8_t level_bit(uint64_t x, uint32_t num_bits, uint32_t i) {
    return (uint8_t)((x >> (num_bits - 1 - i)) & 1);
}

/* ---- dpf.ClientInitialize + GenerateTwoServer (dpf/client.go:17-150) --------------------------- */
void orc_gen_two_server(const uint8_t tape[ORC_TAPE_BYTES], uint32_t num_bits, uint64_t a,
                        uint64_t b, pir_query* q0, pir_query* q1) {
    pir_query* q[2] = {q0, q1};
    memset(q0, 0, sizeof *q0);
    memset(q1, 0, sizeof *q1);
    /* ClientInitialize: four rand.Read(16) PRF keys (dpf/client.go:23-35) */
    for (int p = 0; p < 2; p++) {
        memcpy(q[p]->prf_keys, tape, 64);
        q[p]->num_bits = num_bits;
        q[p]->share_number = (uint8_t)p;
        q[p]->is_two_party = 1;
        q[p]->group_size = 1;
    }
    uint8_t rk[3][176];
    for (int i = 0; i < 3; i++) orc_aes128_expand_key(tape + 16 * i, rk[i]);
    /* dpf/client.go:59-68 */
    const uint8_t* r1 = tape + 64;
    memcpy(q0->s_init, r1, 16);
    q0->t_init = r1[16] % 2;
    memcpy(q1->s_init, tape + 81, 16);
    q1->t_init = q0->t_init ^ 1;

    uint8_t s0[16], s1[16];
    memcpy(s0, q0->s_init, 16);
    memcpy(s1, q1->s_init, 16);
    uint8_t t0 = q0->t_init, t1 = q1->t_init;
    const int leftStart = 0, rightStart = 17;
    for (uint32_t i = 0; i < num_bits; i++) {
        uint8_t o0[48], o1[48];
        orc_mmo3(rk, s0, o0); /* dpf/client.go:91-96 */
        orc_mmo3(rk, s1, o1);
        uint8_t t0L = o0[16] % 2, t0R = o0[33] % 2, t1L = o1[16] % 2, t1R = o1[33] % 2;
        uint8_t aBit = level_bit(a, num_bits, i);
        int keep = rightStart, lose = leftStart;
        if (aBit == 0) {
            keep = leftStart;
            lose = rightStart;
        }
        uint8_t* cw = q0->cw[i];
        for (int j = 0; j < 16; j++) cw[j] = o0[lose + j] ^ o1[lose + j]; /* :118-121 */
        cw[16] = t0L ^ t1L ^ aBit ^ 1;
        cw[17] = t0R ^ t1R ^ aBit;
        for (int j = 0; j < 16; j++) { /* :127-130 */
            s0[j] = o0[keep + j] ^ (uint8_t)(t0 * cw[j]);
            s1[j] = o1[keep + j] ^ (uint8_t)(t1 * cw[j]);
        }
        uint8_t tCWKeep = (keep == rightStart) ? cw[17] : cw[16];
        t0 = (uint8_t)((o0[keep + 16] % 2) ^ (uint8_t)(tCWKeep * t0)); /* :137-138 */
        t1 = (uint8_t)((o1[keep + 16] % 2) ^ (uint8_t)(tCWKeep * t1));
        memcpy(q1->cw[i], cw, PIR_CW_BYTES);
    }
    /* dpf/client.go:141-148, Go int arithmetic wraps */
    uint64_t f = (uint64_t)b - (uint64_t)orc_varint8(s0) + (uint64_t)orc_varint8(s1);
    if (t1 == 1) f = (uint64_t)0 - f;
    q0->final_cw = (int64_t)f;
    q1->final_cw = (int64_t)f;
}

/* ---- dpf.ServerInitialize (dpf/server.go:16-46): AES key schedules of the shipped keys ---------- */
void orc_server_initialize(const pir_query* q, uint32_t num_bits, orc_dpf* f) {
    for (int i = 0; i < 3; i++) orc_aes128_expand_key(q->prf_keys[i], f->rk[i]);
    f->num_bits = num_bits;
}

/* ---- (*Dpf).Evaluate2P (dpf/server.go:55-101) -------------------------------------------------- */
int64_t orc_evaluate2p(const orc_dpf* f, const pir_query* q, uint64_t x) {
    uint8_t s[16], out[48];
    memcpy(s, q->s_init, 16);
    uint8_t t = q->t_init;
    for (uint32_t i = 0; i < f->num_bits; i++) {
        uint8_t xBit = level_bit(x, f->num_bits, i);
        orc_mmo3(f->rk, s, out);
        /* G(s) ^ t*(sCW||tLCW||sCW||tRCW) over bytes 0..33 (dpf/server.go:72-83) */
        const uint8_t* cw = q->cw[i];
        int count = 0;
        for (int j = 0; j < 34; j++) {
            if (j == 17)
                count = 0;
            else if (j == 33)
                count = 17;
            out[j] = out[j] ^ (uint8_t)(t * cw[count]);
            count++;
        }
        if (xBit == 0) {
            memcpy(s, out, 16);
            t = out[16] % 2;
        } else {
            memcpy(s, out + 17, 16);
            t = out[33] % 2;
        }
    }
    uint64_t v = (uint64_t)orc_varint8(s) + (uint64_t)t * (uint64_t)q->final_cw;
    if (q->share_number == 0) return (int64_t)v;
    return (int64_t)((uint64_t)0 - v);
}

/* db.go:114-121: dimHeight = DBSize/GroupSize (integer), numBits = uint(log2(dimHeight)+1) */
uint32_t orc_num_bits_for(uint64_t db_size, uint32_t group_size, int is_keyword) {
    if (is_keyword) return 32;
    uint64_t h = db_size / group_size;
    uint32_t nb = 0;
    while (h) {
        nb++;
        h >>= 1;
    }
    return nb; /* floor(log2 h) + 1 */
}

/* bits[i] = (|res % 2| == 0) (db.go:142) */
static inline uint8_t even_bit(int64_t res) { return (uint8_t)((res % 2) == 0); }

typedef struct expand_job {
    const pir_query* q;
    const orc_dpf* f;
    const uint64_t* keywords;
    uint64_t row_begin, n_rows;
    uint8_t* bits;
    /* tree variant */
    uint64_t height;
} expand_job;

static void* expand_rows_worker(void* p) {
    expand_job* j = (expand_job*)p;
    for (uint64_t r = 0; r < j->n_rows; r++) {
        uint64_t row = j->row_begin + r;
        uint64_t key = j->q->is_keyword ? j->keywords[row] : row; /* db.go:132-135 */
        j->bits[r] = even_bit(orc_evaluate2p(j->f, j->q, key));
    }
    return NULL;
}

static void run_split(expand_job* base, int nthreads, void* (*fn)(void*), uint64_t grain) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads == 1 || base->n_rows < 2 * grain) {
        fn(base);
        return;
    }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    expand_job* jobs = (expand_job*)malloc(sizeof(expand_job) * (size_t)nthreads);
    uint64_t chunks = (base->n_rows + grain - 1) / grain;
    uint64_t per = (chunks + (uint64_t)nthreads - 1) / (uint64_t)nthreads;
    int started = 0;
    for (int t = 0; t < nthreads; t++) {
        uint64_t b = (uint64_t)t * per * grain;
        if (b >= base->n_rows) break;
        uint64_t e = b + per * grain;
        if (e > base->n_rows) e = base->n_rows;
        jobs[t] = *base;
        jobs[t].row_begin = base->row_begin + b;
        jobs[t].n_rows = e - b;
        jobs[t].bits = base->bits + b;
        pthread_create(&th[t], NULL, fn, &jobs[t]);
        started++;
    }
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

/* ---- (*Database).ExpandSharedQuery (db.go:110-174) ---------------------------------------------- */
void orc_expand_shared_query(const pir_query* q, uint64_t db_size, const uint64_t* keywords,
                             uint64_t row_begin, uint64_t n_rows, uint8_t* bits, int nthreads) {
    orc_dpf f;
    orc_server_initialize(q, orc_num_bits_for(db_size, q->group_size, q->is_keyword), &f);
    expand_job j = {q, &f, keywords, row_begin, n_rows, bits, 0};
    run_split(&j, nthreads, expand_rows_worker, 64);
}

/* ---- the same bits by GGM tree expansion ------------------------------------------------------- */
static void tree_rec(const expand_job* j, const uint8_t s[16], uint8_t t, uint32_t level, uint64_t prefix,
                     uint64_t lo, uint64_t hi) {
    uint32_t nb = j->f->num_bits;
    if (level == nb) {
        uint64_t v = (uint64_t)orc_varint8(s) + (uint64_t)t * (uint64_t)j->q->final_cw;
        if (j->q->share_number != 0) v = (uint64_t)0 - v;
        j->bits[prefix - j->row_begin] = even_bit((int64_t)v);
        return;
    }
    uint8_t out[48];
    orc_mmo3(j->f->rk, s, out);
    const uint8_t* cw = j->q->cw[level];
    if (t) {
        for (int k = 0; k < 16; k++) {
            out[k] ^= cw[k];
            out[17 + k] ^= cw[k];
        }
        out[16] ^= cw[16];
        out[33] ^= cw[17];
    }
    uint32_t rem = nb - level - 1; /* bits below the children */
    for (uint64_t c = 0; c < 2; c++) {
        uint64_t child = (prefix << 1) | c;
        uint64_t first = child << rem, last = first + ((uint64_t)1 << rem); /* rows [first,last) */
        if (first >= hi || last <= lo) continue;
        if (c == 0)
            tree_rec(j, out, out[16] & 1, level + 1, child, lo, hi);
        else
            tree_rec(j, out + 17, out[33] & 1, level + 1, child, lo, hi);
    }
}

static void* expand_tree_worker(void* p) {
    expand_job* j = (expand_job*)p;
    tree_rec(j, j->q->s_init, j->q->t_init, 0, 0, j->row_begin, j->row_begin + j->n_rows);
    return NULL;
}

void orc_expand_tree(const pir_query* q, uint64_t db_size, uint64_t row_begin, uint64_t n_rows,
                     uint8_t* bits, int nthreads) {
    orc_dpf f;
    orc_server_initialize(q, orc_num_bits_for(db_size, q->group_size, 0), &f);
    expand_job j = {q, &f, NULL, row_begin, n_rows, bits, 0};
    run_split(&j, nthreads, expand_tree_worker, 4096);
}

/* ---- PrivateSecretSharedQueryWithExpandedBits (db.go:75-107), XorSlots (slot.go:19-30) --------- */
void orc_answer_with_bits(const uint8_t* slots, uint32_t slot_bytes, uint32_t group_size,
                          const uint8_t* bits, uint64_t n_rows, uint8_t* out) {
    size_t row_bytes = (size_t)group_size * slot_bytes;
    memset(out, 0, row_bytes);
    for (uint64_t row = 0; row < n_rows; row++) {
        if (!bits[row]) continue;
        const uint8_t* src = slots + row * row_bytes;
        for (size_t k = 0; k < row_bytes; k++) out[k] ^= src[k];
    }
}

void orc_answer(const pir_query* q, const uint8_t* slots, uint64_t db_size, uint32_t slot_bytes,
                const uint64_t* keywords, uint8_t* out, int nthreads, int use_tree) {
    uint64_t h = db_size / q->group_size;
    uint8_t* bits = (uint8_t*)malloc(h ? h : 1);
    if (use_tree && !q->is_keyword)
        orc_expand_tree(q, db_size, 0, h, bits, nthreads);
    else
        orc_expand_shared_query(q, db_size, keywords, 0, h, bits, nthreads);
    orc_answer_with_bits(slots, slot_bytes, q->group_size, bits, h, out);
    free(bits);
}

/* ---- synthetic data ---------------------------------------------------------------------------- */
uint64_t orc_synth_word(uint64_t seed, uint64_t slot, uint32_t w) {
    uint64_t z = seed + slot * 0x9E3779B97F4A7C15ull + (uint64_t)w * 0xD1B54A32D192ED03ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

void orc_synth_fill(uint64_t seed, uint64_t slot_begin, uint64_t n_slots, uint32_t slot_bytes,
                    uint8_t* out) {
    for (uint64_t i = 0; i < n_slots; i++) {
        uint8_t* dst = out + i * slot_bytes;
        for (uint32_t j = 0; j < slot_bytes; j++) {
            uint64_t w = orc_synth_word(seed, slot_begin + i, j / 8);
            dst[j] = (uint8_t)(w >> (8 * (j % 8)));
        }
    }
}

/* ---- answers over the synthetic database without materialising it ------------------------------ */
/* PrivateSecretSharedQueryWithExpandedBits (db.go:75-107) for nq bit vectors at once over the database that
 * orc_synth_fill(seed) defines: every row of rows [row_begin, row_begin + n_rows) is generated on the fly and
 * XORed (slot.go:19-30) into the answer of every query whose bit for it is set.  bits is [nq][n_rows], out is
 * [nq][group_size*slot_bytes] (zeroed here).  Rows are split statically over nthreads.  This is what lets a
 * test or bench.py check a full-size (4-8 GiB) GPU answer against the restated algorithm on the host. */
typedef struct synth_job {
    uint64_t seed, row_begin, n_rows, stride;
    uint32_t slot_bytes, group_size, nq;
    const uint8_t* bits;
    uint8_t* acc; /* [nq][row_bytes], private to the thread */
} synth_job;

static void* synth_answer_worker(void* p) {
    synth_job* j = (synth_job*)p;
    const size_t row_bytes = (size_t)j->group_size * j->slot_bytes;
    uint8_t* row = (uint8_t*)malloc(row_bytes + 8);
    for (uint64_t r = 0; r < j->n_rows; r++) {
        int any = 0;
        for (uint32_t q = 0; q < j->nq && !any; q++) any = j->bits[(size_t)q * j->stride + r] != 0;
        if (!any) continue; /* db.go:93 */
        const uint64_t slot0 = (j->row_begin + r) * j->group_size;
        if ((j->slot_bytes & 7u) == 0) { /* whole words: the same bytes as orc_synth_fill, eight at a time */
            uint64_t* w = (uint64_t*)row;
            const uint32_t wps = j->slot_bytes / 8;
            for (uint32_t c = 0; c < j->group_size; c++)
                for (uint32_t k = 0; k < wps; k++) w[(size_t)c * wps + k] = orc_synth_word(j->seed, slot0 + c, k);
        } else {
            orc_synth_fill(j->seed, slot0, j->group_size, j->slot_bytes, row);
        }
        for (uint32_t q = 0; q < j->nq; q++) {
            if (!j->bits[(size_t)q * j->stride + r]) continue;
            uint8_t* a = j->acc + (size_t)q * row_bytes;
            size_t k = 0;
            for (; k + 8 <= row_bytes; k += 8) { /* XorSlots, eight bytes at a time */
                uint64_t x, y;
                memcpy(&x, a + k, 8);
                memcpy(&y, row + k, 8);
                x ^= y;
                memcpy(a + k, &x, 8);
            }
            for (; k < row_bytes; k++) a[k] ^= row[k];
        }
    }
    free(row);
    return NULL;
}

void orc_answers_over_synth(uint64_t seed, uint32_t slot_bytes, uint32_t group_size, uint64_t row_begin,
                            uint64_t n_rows, const uint8_t* bits, uint32_t nq, uint8_t* out, int nthreads) {
    const size_t row_bytes = (size_t)group_size * slot_bytes;
    memset(out, 0, (size_t)nq * row_bytes);
    if (nthreads < 1) nthreads = 1;
    if ((uint64_t)nthreads > n_rows) nthreads = n_rows ? (int)n_rows : 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    synth_job* jobs = (synth_job*)malloc(sizeof(synth_job) * (size_t)nthreads);
    for (int t = 0; t < nthreads; t++) {
        const uint64_t b = n_rows * (uint64_t)t / (uint64_t)nthreads, e = n_rows * (uint64_t)(t + 1) / (uint64_t)nthreads;
        synth_job j = {seed, row_begin + b, e - b, n_rows, slot_bytes, group_size, nq, bits + b,
                       (uint8_t*)calloc((size_t)nq ? (size_t)nq : 1, row_bytes)};
        jobs[t] = j;
        pthread_create(&th[t], NULL, synth_answer_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) {
        pthread_join(th[t], NULL);
        for (size_t k = 0; k < (size_t)nq * row_bytes; k++) out[k] ^= jobs[t].acc[k];
        free(jobs[t].acc);
    }
    free(th);
    free(jobs);
}

void orc_tape(uint64_t seed, uint8_t* out, size_t n) {
    uint64_t x = seed;
    for (size_t i = 0; i < n; i++) {
        if (i % 8 == 0) {
            x += 0x9E3779B97F4A7C15ull;
        }
        uint64_t z = x;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z = z ^ (z >> 31);
        out[i] = (uint8_t)(z >> (8 * (i % 8)));
    }
}
