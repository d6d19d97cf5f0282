/*
 * pir_b200.h — C ABI of the B200-native server-side answer path for the two-server DPF PIR of
 * mvmcconnell/pir.  This is the drop-in boundary: everything a cgo (or ctypes / C++) binding of the
 * reference's hot path needs, with plain pointers and sizes only.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   dpf.ServerInitialize                      dpf/server.go:16-46   -> key schedule happens inside every call
 *   (*Dpf).Evaluate2P                         dpf/server.go:55-101  -> pir_dpf_eval_points
 *   (*Database).ExpandSharedQuery             db.go:110-174         -> pir_expand_bits
 *   (*Database).PrivateSecretSharedQueryWithExpandedBits db.go:75-107 -> pir_answer_with_bits
 *   (*Database).PrivateSecretSharedQuery      db.go:68-72           -> pir_answer / pir_answer_batch
 *   XorSlots / Recover                        slot.go:19-30, query.go:278-297 -> pir_xor_combine (host)
 *   Database{Slots []*Slot}                   db.go:22-26           -> pir_db handle (flat device copy)
 *   Database.SetKeywords                      db.go:392-394         -> pir_db_set_keywords
 *   QueryShare + dpf.Key2P + dpf.PrfKey       query.go:14-22, dpf/common.go:13-35 -> pir_query POD
 *   GenerateAuditForSharedQuery(WithExpandedBits) aspir.go:246-283   -> pir_audit_shared_query(_with_bits)
 *
 * Conventions
 *   - every function returns PIR_OK (0) or a negative pir_status; pir_last_error() gives a
 *     thread-local human-readable message for the last failure on the calling thread.
 *   - the library never frees or keeps caller memory; outputs go to caller-allocated buffers.
 *   - a pir_db handle is bound to ONE CUDA device and owns a flat copy of (a shard of) the slots.
 *     Every entry point is thread-safe.  Calls on one handle ENQUEUE under an internal mutex; single index queries
 *     (pir_answer, pir_server_answer) then wait for their own completion event with the mutex released, so up to 8
 *     queries of different threads are in flight on a handle and their kernels overlap on the GPU.  Batches, keyword
 *     queries, expand / with-bits calls hold the mutex until they return.  Different handles are independent.
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point fails with
 *     PIR_ERR_CUDA.
 */
#ifndef PIR_B200_H
#define PIR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIR_B200_ABI_VERSION 1
#define PIR_MAX_NUM_BITS 64   /* dpf tests use 64-bit domains (dpf/dpf_test.go:73-107) */
#define PIR_CW_BYTES 18       /* sCW[16] || tLCW || tRCW  (dpf/client.go:83-84,118-125) */
#define PIR_NUM_PRF_KEYS 4    /* initPRFLen (dpf/common.go:10); only keys 0..2 are used by 2-party eval */

typedef enum pir_status {
    PIR_OK = 0,
    PIR_ERR_BAD_ARG = -1,     /* null pointer, zero size, group size not dividing a shard, ... */
    PIR_ERR_BAD_KEY = -2,     /* num_bits out of range, t_init not 0/1, share_number > 1 ... */
    PIR_ERR_CUDA = -3,        /* any CUDA runtime failure (message carries cudaGetErrorString) */
    PIR_ERR_OOM = -4,         /* host or device allocation failed */
    PIR_ERR_UNSUPPORTED = -5  /* multi-party (>2 shares) keys: unreachable in the reference too
                                 (dpf/client.go:152-156 panics "not implemented") */
} pir_status;

/*
 * One server's share of a query: QueryShare (query.go:14-22) with its Key2P (dpf/common.go:30-35)
 * and the four shipped PRF keys (dpf/common.go:13-15) flattened into a POD.  160+ bytes of header
 * followed by 64 fixed correction-word slots, of which the first num_bits are meaningful.  This is
 * also the wire format the reference lacks.
 */
typedef struct pir_query {
    uint8_t  prf_keys[PIR_NUM_PRF_KEYS][16]; /* QueryShare.PrfKeys[i].Bytes                         */
    uint8_t  s_init[16];                     /* Key2P.SInit                                         */
    uint8_t  t_init;                         /* Key2P.TInit, 0 or 1                                 */
    uint8_t  share_number;                   /* QueryShare.ShareNumber, 0 or 1                      */
    uint8_t  is_keyword;                     /* QueryShare.IsKeywordBased                           */
    uint8_t  is_two_party;                   /* QueryShare.IsTwoParty, must be 1                    */
    uint32_t num_bits;                       /* len(Key2P.CW); informational for pir_answer*, which
                                                recompute it like db.go:117-121 and reject a mismatch */
    int64_t  final_cw;                       /* Key2P.FinalCW (Go int, wrapping int64)              */
    uint32_t group_size;                     /* QueryShare.GroupSize                                */
    uint32_t reserved;                       /* must be 0                                           */
    uint8_t  cw[PIR_MAX_NUM_BITS][PIR_CW_BYTES]; /* Key2P.CW[i]                                     */
} pir_query;

typedef struct pir_db pir_db; /* opaque */

/* ---- library / device ------------------------------------------------------------------------ */
int         pir_abi_version(void);
const char* pir_last_error(void);
/* Number of visible CUDA devices (0 when there is none); never fails. */
int         pir_device_count(void);

/* ---- database handle (replaces Database{DBMetadata, Slots, Keywords}, db.go:15-26) ------------ */
/*
 * Create a handle on `device` for the slots [slot_begin, slot_begin+slot_count) of a logical
 * database of db_size slots x slot_bytes bytes.  A single-GPU server passes slot_begin = 0,
 * slot_count = db_size.  A row-sharded multi-GPU server creates one handle per GPU; a query with
 * group size W then needs slot_begin % W == 0 and (slot_count % W == 0 or the shard is the last).
 * Slots are stored flat with a pitch of slot_bytes rounded up to 16 (pir_db_slot_pitch).
 */
int pir_db_create(int device, uint64_t db_size, uint32_t slot_bytes,
                  uint64_t slot_begin, uint64_t slot_count, pir_db** out);
int pir_db_destroy(pir_db* db);
/* Copy n_slots densely packed slots (n_slots*slot_bytes host bytes) into global slots
 * [slot_index, slot_index+n_slots), which must lie inside the shard. */
int pir_db_upload(pir_db* db, uint64_t slot_index, const void* host_slots, uint64_t n_slots);
/* Read slots back (testing / Database round trips). */
int pir_db_download(pir_db* db, uint64_t slot_index, void* host_slots, uint64_t n_slots);
/* Fill the shard on the device with the seeded synthetic generator of DESIGN.md ("pir_synth"):
 * byte j of slot i is a pure function of (seed, i, j), identical to oracle/pir_oracle.c orc_synth_fill. */
int pir_db_fill_synthetic(pir_db* db, uint64_t seed);
/* Keywords for the shard's slots (Database.SetKeywords, db.go:392-394): n must equal slot_count. */
int pir_db_set_keywords(pir_db* db, const uint64_t* keywords, uint64_t n);
/* Keyword queries with group size W > 1 evaluate Keywords[r] at ROW r (db.go:132-135), and on every shard but the first
 * r = slot / W lies below the shard's own slots: row_keywords[k] = Keywords[first_row + k] for the n = n_rows rows that
 * pir_db_rows(db, group_size) reports.  One group size at a time (a new call replaces the last); not needed on an
 * unsharded handle; pir_server_set_keywords calls it on every shard for the server's shard_align. */
int pir_db_set_row_keywords(pir_db* db, uint32_t group_size, const uint64_t* row_keywords, uint64_t n);
uint32_t pir_db_slot_pitch(const pir_db* db);
int pir_db_device(const pir_db* db);
/*
 * Path selection and work-split overrides, for tests and measurements.  The library reads NO environment
 * variable; these options are the only switches it has, and every combination of them gives bit-identical
 * answers (the test-suite runs the paths against each other and against the oracle).  Defaults are the
 * measured optimum.  Names (value range; default first):
 *   dfs_depth (-1 = cost model, 0..7), hi_warps (-1, 4|8|12|16), dmin (-1, 0..7)   work split of the answer kernel
 *   top_warp (-1 = by row size, 0, 1)     the last scan warp walks the tops of the tasks ahead
 *   slice_log (0 = by task depth, 2, 3)   rows per thread per hand-off slice (log2)
 *   batch_mode (0 = by row size, 1 = one fused launch per query, 2 = bit vectors + one scan pass)
 *   batch_one_stream, batch_latency_split, host_copy_path, signal_kernel, serial_launch, no_lookahead, no_trim, direct_combine (0, 1)
 *   kernel_v1, expand_v1, keyword_v1, force_unfused (0, 1)   the older two-launch / per-row kernels
 * PIR_ERR_BAD_ARG for an unknown name or a value out of range.
 */
int pir_db_set_option(pir_db* db, const char* name, int64_t value);

/* ---- the answer path -------------------------------------------------------------------------- */
/*
 * pir_answer: PrivateSecretSharedQuery (db.go:68-72).  out receives group_size*slot_bytes bytes:
 * Shares[c].Data = out[c*slot_bytes .. (c+1)*slot_bytes).  For a shard handle the result is the
 * shard's partial answer; XOR the partials of all shards (pir_xor_combine) for the server's answer.
 */
int pir_answer(pir_db* db, const pir_query* q, void* out, size_t out_len);
/* n_queries independent queries answered in one pass over the database; out is n_queries
 * consecutive answers of group_size*slot_bytes bytes (all queries must share group_size and
 * is_keyword). */
int pir_answer_batch(pir_db* db, const pir_query* qs, uint32_t n_queries, void* out, size_t out_len);
/* ExpandSharedQuery (db.go:110-174): bits_out[r] in {0,1} for the shard's rows, one byte per row
 * like Go's []bool.  n_rows must be the shard's row count for q->group_size. */
int pir_expand_bits(pir_db* db, const pir_query* q, uint8_t* bits_out, uint64_t n_rows);
/*
 * pir_expand_bits without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same kernels as pir_expand_bits, with their scratch (the packed selection-bit vector) in a file-scope static
 * device array of PIR_NOALLOC_MAX_ROWS bits (2 MiB, held by every GPU that loads the library until exit) and the result
 * -- one 0/1 byte per shard row -- written to the CALLER's device buffer d_bits_out (n_rows bytes, 4-byte aligned).
 * Only enqueues work on `stream` (a cudaStream_t as void*, NULL = the handle's own): no cudaMalloc / cudaFree, no
 * synchronisation.  Index queries only; n_rows must be the shard's row count and at most PIR_NOALLOC_MAX_ROWS
 * (PIR_ERR_BAD_ARG beyond it).  ALL calls on one GPU share the array: each call clears what it uses first, and calls
 * whose GPU work overlaps (different streams, different handles on the same GPU) race -- serialise them.
 */
#define PIR_NOALLOC_MAX_ROWS (1ull << 24)
int pir_expand_bits_noalloc(pir_db* db, const pir_query* q, uint8_t* d_bits_out, uint64_t n_rows, void* stream);
/* PrivateSecretSharedQueryWithExpandedBits (db.go:75-107). */
int pir_answer_with_bits(pir_db* db, uint32_t group_size, const uint8_t* bits, uint64_t n_rows,
                         void* out, size_t out_len);
/* Rows of the shard for a given group size: floor(db_size/W) restricted to the shard. */
int pir_db_rows(const pir_db* db, uint32_t group_size, uint64_t* first_row, uint64_t* n_rows);

/*
 * Device-resident variants for callers that keep queries/answers in GPU memory and want to time
 * the kernels alone: d_out is a device pointer (group_size*slot_bytes*n_queries bytes), stream is a
 * cudaStream_t passed as void* (NULL = the handle's own stream).  The call only enqueues work.
 */
int pir_answer_batch_device(pir_db* db, const pir_query* qs, uint32_t n_queries, void* d_out,
                            void* stream);

/* ---- multi-GPU: one PIR server row-sharded over the GPUs of a box ---------------- */
/*
 * The answer is an XOR over rows (db.go:91-104), so GPU g holds rows [g*H/G, (g+1)*H/G) and every
 * shard's kernel XORs its partial answer straight into ONE combine slot with red.global.xor — on the
 * owner GPU that is local memory, on the others it is the owner's memory mapped over NVLink (peer
 * access or CUDA IPC).  No NCCL (it has no XOR reduction), no extra combine pass.
 *
 * pir_server: ONE process drives all shards (what a Go server does: one cgo call per query).
 */
typedef struct pir_server pir_server; /* opaque */
/* devices[i] is the CUDA device of shard i (a device may appear more than once: several shards on one
 * GPU).  Shard boundaries are multiples of shard_align slots (pass the largest group size the
 * database will be queried with, or 1). */
int pir_server_create(const int* devices, int n_devices, uint64_t db_size, uint32_t slot_bytes,
                      uint32_t shard_align, pir_server** out);
int pir_server_destroy(pir_server* s);
int pir_server_num_shards(const pir_server* s);
/* Borrow shard i's handle (owned by the server) for upload / keywords / per-shard calls. */
int pir_server_shard(pir_server* s, int i, pir_db** db);
int pir_server_upload(pir_server* s, uint64_t slot_index, const void* host_slots, uint64_t n_slots);
int pir_server_fill_synthetic(pir_server* s, uint64_t seed);
/* Database.SetKeywords (db.go:392-394) for the whole database: n must equal db_size; shard g keeps
 * keywords[slot_begin_g, +slot_count_g) and, when shard_align > 1, the keywords of its rows at that group size
 * (pir_db_set_row_keywords).  Keyword queries (db.go:119-121,132-135) take group_size 1 or shard_align. */
int pir_server_set_keywords(pir_server* s, const uint64_t* keywords, uint64_t n);
/* pir_db_set_option on every shard. */
int pir_server_set_option(pir_server* s, const char* name, int64_t value);
/* PrivateSecretSharedQuery (db.go:68-72) over all shards: out gets group_size*slot_bytes bytes. */
int pir_server_answer(pir_server* s, const pir_query* q, void* out, size_t out_len);
int pir_server_answer_batch(pir_server* s, const pir_query* qs, uint32_t n_queries, void* out, size_t out_len);

/*
 * pir_combine: the same combine for ONE PROCESS PER GPU (torchrun-style launch).  The owner rank
 * creates a ring of n_slots answer slots of slot_stride bytes plus one arrival counter per slot and
 * exports a CUDA IPC handle; every other rank opens it.  Per query every rank calls
 * pir_answer_accumulate_device on the same slot; the owner then waits for `world` arrivals and reads
 * the combined answer.  A slot must be re-zeroed by the owner (pir_combine_reset) before it is reused,
 * and no rank may accumulate into it before that reset is complete (the caller's barrier).
 */
#define PIR_IPC_HANDLE_BYTES 64
typedef struct pir_combine pir_combine; /* opaque */
int pir_combine_create(int device, uint32_t n_slots, size_t slot_stride, pir_combine** out);
int pir_combine_export(pir_combine* c, uint8_t handle[PIR_IPC_HANDLE_BYTES]);
int pir_combine_open(int device, const uint8_t handle[PIR_IPC_HANDLE_BYTES], uint32_t n_slots,
                     size_t slot_stride, pir_combine** out);
int pir_combine_destroy(pir_combine* c);
/* Owner only: zero slots [first, first+count) and their arrival counters; blocks until done. */
int pir_combine_reset(pir_combine* c, uint32_t first, uint32_t count);
/* Any rank: XOR this shard's partial answer of q into slot `slot` (pitched layout, like
 * pir_answer_batch_device) and signal arrival.  Only enqueues work on `stream` (NULL = the handle's). */
int pir_answer_accumulate_device(pir_db* db, const pir_query* q, pir_combine* c, uint32_t slot, void* stream);
/* Owner only: wait (on the device, bounded by timeout_ms) until `expected` ranks have arrived at the
 * slot, then copy its first out_len bytes to host memory.  PIR_ERR_CUDA on timeout. */
int pir_combine_wait_read(pir_combine* c, uint32_t slot, uint32_t expected, void* out, size_t out_len,
                          uint32_t timeout_ms);
/* Owner only: the same wait, then the slot and its arrival counter are zeroed behind the copy (one synchronisation):
 * on return the slot is free for its next query. */
int pir_combine_wait_read_reset(pir_combine* c, uint32_t slot, uint32_t expected, void* out, size_t out_len,
                                uint32_t timeout_ms);
/* Owner only: enqueue the (device-side, bounded) wait for `expected` arrivals on `stream` (NULL = the combine's own)
 * and return; work enqueued on that stream afterwards sees the combined answer in the slot. */
int pir_combine_wait_device(pir_combine* c, uint32_t slot, uint32_t expected, uint32_t timeout_ms, void* stream);
/* Device address of a slot in this process (for device-resident consumers / tests). */
void* pir_combine_slot_ptr(pir_combine* c, uint32_t slot);

/* ---- secret-shared ASPIR audit (aspir.go:227-295) ------------------------------ */
/*
 * GenerateAuditForSharedQuery (aspir.go:246-257): the key database is queried with group size 1 whatever
 * the query says (aspir.go:251-254), and the server's auth-token share is XORed into the result
 * (aspir.go:279-281, XorSlots semantics: min of the two lengths).  out receives slot_bytes bytes = the
 * AuditTokenShare.  Like the reference (aspir.go:275-277 after db.go:79-89), a query whose own group size
 * is not 1 is rejected with PIR_ERR_BAD_ARG ("Invalid challenge ciphertext result").
 */
int pir_audit_shared_query(pir_db* key_db, const pir_query* q, const void* auth_token, size_t token_len,
                           void* out, size_t out_len);
/* GenerateAuditForSharedQueryWithExpandedBits (aspir.go:261-283): the bits of ONE expansion can serve
 * the data database and the key database. */
int pir_audit_shared_query_with_bits(pir_db* key_db, uint32_t group_size, const uint8_t* bits, uint64_t n_rows,
                                     const void* auth_token, size_t token_len, void* out, size_t out_len);

/* ---- dpf package surface ---------------------------------------------------------------------- */
/* (*Dpf).Evaluate2P (dpf/server.go:55-101) on n points: out[i] = Evaluate2P(share_number, key, xs[i])
 * as wrapping int64.  Uses q->num_bits levels; runs on `device`. */
int pir_dpf_eval_points(int device, const pir_query* q, const uint64_t* xs, uint64_t n, int64_t* out);

/* ---- client side (host only; the "next" row N1 of SURVEY.md 8f) ------------------------------- */
/* dpf.ClientInitialize + (*Dpf).GenerateTwoServer(a, b) (dpf/client.go:17-150): fills both servers'
 * keys (prf keys, s_init, t_init, cw, final_cw, share_number, num_bits; group_size = 1, is_keyword = 0
 * — the caller sets those two like newQueryShares does, query.go:83-98).  tape = the 97 bytes the
 * reference would draw from crypto/rand in its order (4x16 PRF keys, 17, 16), or NULL for /dev/urandom. */
int pir_client_gen_two_server(uint32_t num_bits, uint64_t a, uint64_t b, const uint8_t* tape,
                              pir_query* q0, pir_query* q1);

/* ---- wire encoding of a query share (host only) ------------------------------------------------ */
/* The reference has no serialisation for QueryShare / Key2P (query.go:14-22, dpf/common.go:30-35 are in-process structs),
 * yet the two servers receive their shares over a network.  pir_query above is the in-memory form (1336 bytes, host
 * endianness); this is the compact, endianness-defined form: 100 + 18*num_bits bytes, every integer little-endian:
 *   0  "PIRQ"            4   magic
 *   4  version           1   PIR_WIRE_VERSION
 *   5  flags             1   bit0 share_number, bit1 is_keyword, bit2 t_init, bit3 is_two_party; bits 4-7 zero
 *   6  num_bits          1   1..PIR_MAX_NUM_BITS
 *   7  zero              1
 *   8  group_size        4
 *  12  final_cw          8   two's complement
 *  20  s_init           16
 *  36  prf_keys         64   4 x 16 (dpf/common.go:10 ships four)
 * 100  cw               18 * num_bits
 * decode rejects anything else (wrong magic/version/length, non-zero padding, num_bits out of range) with PIR_ERR_BAD_KEY
 * and leaves *q zeroed; a decoded share is exactly what encode was given (fields beyond num_bits in cw are zero). */
#define PIR_WIRE_VERSION 1
#define PIR_WIRE_HEADER_BYTES 100
size_t pir_query_wire_size(uint32_t num_bits);  /* 0 if num_bits is out of range */
int    pir_query_encode(const pir_query* q, void* buf, size_t buf_len, size_t* written);
int    pir_query_decode(const void* buf, size_t len, pir_query* q);

/* ---- host helpers ----------------------------------------------------------------------------- */
/* acc[i] ^= part[i]: XorSlots (slot.go:19-30) / Recover (query.go:278-297) / multi-GPU combine. */
void pir_xor_combine(void* acc, const void* part, size_t len);
/* numBits as the server computes it (db.go:114-121): floor(log2(rows))+1, or 32 for keyword queries. */
uint32_t pir_num_bits_for(uint64_t db_size, uint32_t group_size, int is_keyword);

/* Per-launch timing of the dominant (fused expand+scan) kernel with CUDA events on the launching
 * stream, for the roofline line of bench.py.  get() waits for the recorded launches, returns their
 * summed duration and count, and resets the recorder. */
int pir_db_set_kernel_timing(pir_db* db, int enable);
int pir_db_get_kernel_timing(pir_db* db, double* total_ms, uint64_t* launches);

/* The AES-only rate of a shard, measured: average duration of `reps` launches of the answer kernel in BITS mode (the
 * same tree expansion with the same work split, no database row read), for bench.py's lookup-pipe roofline. */
int pir_db_time_expansion(pir_db* db, const pir_query* q, uint32_t reps, double* ms_per_launch);

/* counters for benchmarking: number of kernel launches issued by this library in this process */
uint64_t pir_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PIR_B200_H */
