// capi.cu — the C ABI of include/pir_b200.h: handles, key marshalling, launch planning.
// No CPU fallback exists: without a CUDA device every compute entry point returns PIR_ERR_CUDA.
#include "../../include/pir_b200.h"
#include "kernels.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <initializer_list>
#include <mutex>
#include <new>
#include <vector>

#ifdef PIRB_DEV
#include "dev_hooks.cuh"
using pirb_dev::DevPhase;
static inline bool dev_phase_debug_on() { return pirb_dev::dev_phase_debug(); }
#else
static inline bool dev_phase_debug_on() { return false; }
#endif

using namespace pirb;

// ---- errors -------------------------------------------------------------------------------------
static thread_local char tl_err[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tl_err, sizeof tl_err, fmt, ap);
    va_end(ap);
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(e_ == cudaErrorMemoryAllocation ? PIR_ERR_OOM : PIR_ERR_CUDA, "%s: %s",    \
                        #call, cudaGetErrorString(e_));                                            \
    } while (0)

extern "C" const char* pir_last_error(void) { return tl_err; }
extern "C" int pir_abi_version(void) { return PIR_B200_ABI_VERSION; }
extern "C" int pir_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
extern "C" uint64_t pir_kernel_launch_count(void) { return launch_count(); }

// ---- host AES-128 key schedule (FIPS-197 5.2); the S-box is derived, not tabulated --------------
namespace {
struct AesHost {
    uint8_t sbox[256];
    uint32_t t0[256];
    AesHost() {
        auto mul = [](uint8_t a, uint8_t b) {
            uint8_t p = 0;
            while (b) {
                if (b & 1) p ^= a;
                a = (uint8_t)((a << 1) ^ ((a & 0x80) ? 0x1b : 0));
                b >>= 1;
            }
            return p;
        };
        // walk the multiplicative group with generator 3: p = 3^k, q = 3^-k
        uint8_t p = 1, q = 1;
        do {
            p = (uint8_t)(p ^ (p << 1) ^ ((p & 0x80) ? 0x1b : 0));  // p *= 3
            q ^= (uint8_t)(q << 1);                                  // q /= 3
            q ^= (uint8_t)(q << 2);
            q ^= (uint8_t)(q << 4);
            if (q & 0x80) q ^= 0x09;
            uint8_t x = (uint8_t)(q ^ (uint8_t)((q << 1) | (q >> 7)) ^ (uint8_t)((q << 2) | (q >> 6)) ^
                                  (uint8_t)((q << 3) | (q >> 5)) ^ (uint8_t)((q << 4) | (q >> 4)));
            sbox[p] = (uint8_t)(x ^ 0x63);
        } while (p != 1);
        sbox[0] = 0x63;
        for (int x = 0; x < 256; x++) {
            uint8_t s = sbox[x];
            t0[x] = (uint32_t)mul(s, 2) | ((uint32_t)s << 8) | ((uint32_t)s << 16) | ((uint32_t)mul(s, 3) << 24);
        }
    }
    void expand(const uint8_t key[16], uint32_t rk[44]) const {
        uint8_t w[176];
        memcpy(w, key, 16);
        uint8_t rc = 1;
        for (int i = 16; i < 176; i += 4) {
            uint8_t t[4] = {w[i - 4], w[i - 3], w[i - 2], w[i - 1]};
            if ((i & 15) == 0) {
                uint8_t t0_ = t[0];
                t[0] = (uint8_t)(sbox[t[1]] ^ rc);
                t[1] = sbox[t[2]];
                t[2] = sbox[t[3]];
                t[3] = sbox[t0_];
                rc = (uint8_t)((rc << 1) ^ ((rc & 0x80) ? 0x1b : 0));
            }
            for (int j = 0; j < 4; j++) w[i + j] = (uint8_t)(w[i - 16 + j] ^ t[j]);
        }
        for (int i = 0; i < 44; i++)
            rk[i] = (uint32_t)w[4 * i] | ((uint32_t)w[4 * i + 1] << 8) | ((uint32_t)w[4 * i + 2] << 16) |
                    ((uint32_t)w[4 * i + 3] << 24);
    }
};
const AesHost& aes_host() {
    static AesHost a;
    return a;
}

// ---- scratch of pir_expand_bits_noalloc --------------------------------------------------------------
// The packed selection-bit vector that the answer kernel's BITS mode ORs into, as a file-scope device array, so that the
// entry point needs no cudaMalloc.  2 MiB + 16 bytes in the memory of EVERY GPU that loads this library, until exit.
// In host code `g_noalloc_bits` is only a stand-in for the symbol, NOT a device address: its device address is taken
// once per device with cudaGetSymbolAddress (device_ctx below, that device current) and kept in DeviceCtx.
constexpr uint64_t kNoAllocMaxRows = 1ull << 24;                     // PIR_NOALLOC_MAX_ROWS in include/pir_b200.h
constexpr uint64_t kNoAllocWords = kNoAllocMaxRows / 32 + 4;         // ceil(rows / 32) + 1 spill word, rounded up to 4
static __device__ __align__(256) uint32_t g_noalloc_bits[kNoAllocWords];

// ---- per-device context -------------------------------------------------------------------------
struct DeviceCtx {
    bool ready = false;
    uint32_t* noalloc_bits = nullptr;  // device address of g_noalloc_bits on this device
    uint32_t* d_t0 = nullptr;
    uint32_t smem_base = 0;
    int num_sms = 0;
    int clock_khz = 0;
};
std::mutex g_dev_mu;
DeviceCtx g_dev[64];

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int device_ctx(int device, DeviceCtx** out) {
    int n = pir_device_count();
    if (n <= 0) return fail(PIR_ERR_CUDA, "no CUDA device is visible (this library has no CPU path)");
    if (device < 0 || device >= n || device >= 64) return fail(PIR_ERR_BAD_ARG, "device %d out of range (%d visible)", device, n);
    std::lock_guard<std::mutex> lk(g_dev_mu);
    DeviceCtx& c = g_dev[device];
    if (!c.ready) {
        DeviceGuard guard(device);  // (the caller's current device is restored on every path out of this block)
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, device));
        if (prop.major < 10)
            return fail(PIR_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                        prop.major, prop.minor);
        c.num_sms = prop.multiProcessorCount;
        // queried once: cudaDeviceGetAttribute(cudaDevAttrClockRate) takes milliseconds on this driver
        CU(cudaDeviceGetAttribute(&c.clock_khz, cudaDevAttrClockRate, device));
        CU(cudaMalloc(&c.d_t0, 1024));
        CU(cudaMemcpy(c.d_t0, aes_host().t0, 1024, cudaMemcpyHostToDevice));
        CU(cudaStreamSynchronize(0));  // small pageable copies return once staged; the kernels run on non-blocking streams
        CU(configure_kernels());
        CU(probe_smem_base(&c.smem_base));
        if (c.smem_base >= 8192) return fail(PIR_ERR_CUDA, "unexpected dynamic shared memory base %u", c.smem_base);
        CU(cudaGetSymbolAddress((void**)&c.noalloc_bits, g_noalloc_bits));  // (this device is current: DeviceGuard above)
        c.ready = true;
    }
    *out = &c;
    return PIR_OK;
}

uint32_t floor_log2_plus1(uint64_t h) {
    uint32_t nb = 0;
    while (h) {
        nb++;
        h >>= 1;
    }
    return nb;
}
}  // namespace

extern "C" uint32_t pir_num_bits_for(uint64_t db_size, uint32_t group_size, int is_keyword) {
    if (is_keyword) return 32;  // db.go:119-121
    if (group_size == 0) return 0;
    return floor_log2_plus1(db_size / group_size);  // db.go:114-117
}

extern "C" void pir_xor_combine(void* acc, const void* part, size_t len) {
    uint8_t* a = (uint8_t*)acc;
    const uint8_t* b = (const uint8_t*)part;
    for (size_t i = 0; i < len; i++) a[i] ^= b[i];
}

// ---- wire encoding of a query share (layout in include/pir_b200.h) --------------------------------
extern "C" size_t pir_query_wire_size(uint32_t num_bits) {
    if (num_bits == 0 || num_bits > PIR_MAX_NUM_BITS) return 0;
    return (size_t)PIR_WIRE_HEADER_BYTES + (size_t)PIR_CW_BYTES * num_bits;
}

extern "C" int pir_query_encode(const pir_query* q, void* buf, size_t buf_len, size_t* written) {
    if (written) *written = 0;
    if (!q || !buf) return fail(PIR_ERR_BAD_ARG, "null argument");
    const size_t need = pir_query_wire_size(q->num_bits);
    if (!need) return fail(PIR_ERR_BAD_KEY, "num_bits %u out of range", q->num_bits);
    if (q->t_init > 1 || q->share_number > 1 || q->is_keyword > 1 || q->is_two_party > 1 || q->reserved)
        return fail(PIR_ERR_BAD_KEY, "flag fields must be 0 or 1 and reserved must be 0");
    if (buf_len < need) return fail(PIR_ERR_BAD_ARG, "buffer of %zu bytes, %zu needed", buf_len, need);
    uint8_t* b = (uint8_t*)buf;
    memcpy(b, "PIRQ", 4);
    b[4] = PIR_WIRE_VERSION;
    b[5] = (uint8_t)(q->share_number | (q->is_keyword << 1) | (q->t_init << 2) | (q->is_two_party << 3));
    b[6] = (uint8_t)q->num_bits;
    b[7] = 0;
    for (int k = 0; k < 4; k++) b[8 + k] = (uint8_t)(q->group_size >> (8 * k));
    const uint64_t f = (uint64_t)q->final_cw;
    for (int k = 0; k < 8; k++) b[12 + k] = (uint8_t)(f >> (8 * k));
    memcpy(b + 20, q->s_init, 16);
    memcpy(b + 36, q->prf_keys, 64);
    memcpy(b + PIR_WIRE_HEADER_BYTES, q->cw, (size_t)PIR_CW_BYTES * q->num_bits);
    if (written) *written = need;
    return PIR_OK;
}

extern "C" int pir_query_decode(const void* buf, size_t len, pir_query* q) {
    if (!q) return fail(PIR_ERR_BAD_ARG, "null argument");
    memset(q, 0, sizeof(*q));
    if (!buf) return fail(PIR_ERR_BAD_ARG, "null argument");
    const uint8_t* b = (const uint8_t*)buf;
    if (len < PIR_WIRE_HEADER_BYTES) return fail(PIR_ERR_BAD_KEY, "encoded share of %zu bytes is shorter than its header", len);
    if (memcmp(b, "PIRQ", 4) != 0) return fail(PIR_ERR_BAD_KEY, "not an encoded query share (magic)");
    if (b[4] != PIR_WIRE_VERSION) return fail(PIR_ERR_BAD_KEY, "wire version %u, this library reads %u", b[4], PIR_WIRE_VERSION);
    if ((b[5] & 0xF0) || b[7]) return fail(PIR_ERR_BAD_KEY, "non-zero padding in an encoded share");
    const size_t need = pir_query_wire_size(b[6]);
    if (!need) return fail(PIR_ERR_BAD_KEY, "num_bits %u out of range", b[6]);
    if (len != need) return fail(PIR_ERR_BAD_KEY, "encoded share of %zu bytes, %zu expected for num_bits %u", len, need, b[6]);
    q->share_number = b[5] & 1;
    q->is_keyword = (b[5] >> 1) & 1;
    q->t_init = (b[5] >> 2) & 1;
    q->is_two_party = (b[5] >> 3) & 1;
    q->num_bits = b[6];
    uint64_t f = 0;
    for (int k = 0; k < 4; k++) q->group_size |= (uint32_t)b[8 + k] << (8 * k);
    for (int k = 0; k < 8; k++) f |= (uint64_t)b[12 + k] << (8 * k);
    q->final_cw = (int64_t)f;
    memcpy(q->s_init, b + 20, 16);
    memcpy(q->prf_keys, b + 36, 64);
    memcpy(q->cw, b + PIR_WIRE_HEADER_BYTES, (size_t)PIR_CW_BYTES * q->num_bits);
    return PIR_OK;
}

// ---- query marshalling --------------------------------------------------------------------------
static int build_query_const(const pir_query* q, uint32_t num_bits, QueryConst* qc) {
    if (!q) return fail(PIR_ERR_BAD_ARG, "query is null");
    if (!q->is_two_party) return fail(PIR_ERR_UNSUPPORTED, "multi-party keys are not supported (dpf/client.go:152-156)");
    if (q->t_init > 1) return fail(PIR_ERR_BAD_KEY, "t_init must be 0 or 1");
    if (q->share_number > 1) return fail(PIR_ERR_BAD_KEY, "share_number must be 0 or 1");
    if (num_bits == 0 || num_bits > PIR_MAX_NUM_BITS) return fail(PIR_ERR_BAD_KEY, "num_bits %u out of range", num_bits);
    memset(qc, 0, sizeof *qc);
    for (int k = 0; k < 3; k++) aes_host().expand(q->prf_keys[k], qc->rk[k]);
    for (uint32_t i = 0; i < num_bits; i++) {
        const uint8_t* cw = q->cw[i];
        for (int w = 0; w < 4; w++)
            qc->cw[i][w] = (uint32_t)cw[4 * w] | ((uint32_t)cw[4 * w + 1] << 8) | ((uint32_t)cw[4 * w + 2] << 16) |
                           ((uint32_t)cw[4 * w + 3] << 24);
        qc->cw_tl |= (uint64_t)(cw[16] & 1) << i;
        qc->cw_tr |= (uint64_t)(cw[17] & 1) << i;
    }
    for (int w = 0; w < 4; w++)
        qc->s_init[w] = (uint32_t)q->s_init[4 * w] | ((uint32_t)q->s_init[4 * w + 1] << 8) |
                        ((uint32_t)q->s_init[4 * w + 2] << 16) | ((uint32_t)q->s_init[4 * w + 3] << 24);
    qc->t_init = q->t_init;
    qc->fcw_parity = (uint32_t)((uint64_t)q->final_cw & 1u);
    qc->num_bits = num_bits;
    qc->share_number = q->share_number;
    qc->final_cw = q->final_cw;
    return PIR_OK;
}

// ---- database handle ----------------------------------------------------------------------------
// Path selection and work-split overrides (pir_db_set_option).  Every combination gives bit-identical answers;
// the defaults are the measured optimum.  The library reads no environment variable.
struct Tuning {
    int dfs_depth = -1;   // uniform / binary split: depth of the deepest per-thread subtree (0..7), -1 = cost model
    int hi_warps = -1;    // uniform split: 4, 8, 12 or 16 "hi" tree warps, -1 = cost model
    int dmin = -1;        // binary split: smallest task depth, -1 = cost model
    int top_warp = -1;    // 1 / 0: the last scan warp walks the tops of the tasks ahead, -1 = by row size
    int slice_log = 0;    // 2 or 3: log2(rows per thread per slice), 0 = by task depth
    int batch_mode = 0;   // batches: 0 = by row size, 1 = one fused launch per query, 2 = bit vectors + one scan pass
    bool batch_one_stream = false, batch_latency_split = false;
    bool kernel_v1 = false, expand_v1 = false, keyword_v1 = false, force_unfused = false;  // the older kernels
    bool signal_kernel = false;   // multi-GPU: separate arrival kernel instead of the answer kernel's last CTA
    bool host_copy_path = false;  // pir_answer: memset + kernel + copy instead of the one-launch host path
    bool serial_launch = false;   // no programmatic dependent launch between consecutive answer kernels
    bool no_lookahead = false;    // launches do not walk the next launch's tree top (kernels.cuh LaMail)
    bool no_trim = false;         // uniform tasks keep their full shape (147 CTAs on 2^24 rows instead of 148 trimmed ones)
    bool direct_combine = false;  // multi-GPU: every CTA XORs into the (remote) combine slot itself, not only the launch's last CTA
};

constexpr uint32_t kCtaDoneRing = 64;
constexpr uint32_t kHostRing = 8;  // single-query host calls in flight per handle
constexpr size_t kAnswerMaxRowBytes = 4096;  // widest row the warp-specialised kernel takes (answer_kernel_supports)

struct pir_db {
    int device = 0;
    DeviceCtx* ctx = nullptr;
    uint64_t db_size = 0;
    uint32_t slot_bytes = 0;
    uint32_t pitch = 0;
    uint64_t slot_begin = 0, slot_count = 0;
    uint8_t* d_slots = nullptr;
    uint64_t* d_keywords = nullptr;
    // keyword queries with group size W > 1 on a shard: row r evaluates Keywords[r] (db.go:132-135), and r = slot / W lies
    // below the shard's own slots.  pir_db_set_row_keywords keeps the keywords of the shard's ROWS for one W.
    uint64_t* d_row_keywords = nullptr;
    uint32_t rowkw_w = 0;
    uint64_t rowkw_first = 0, rowkw_n = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    Tuning tune;
    // grow-only scratch
    uint4* d_front_s[2] = {nullptr, nullptr};
    uint8_t* d_front_t[2] = {nullptr, nullptr};
    uint64_t front_cap = 0;
    uint8_t* d_bits = nullptr;
    uint64_t bits_cap = 0;
    uint32_t* d_cta_done = nullptr;  // multi-GPU combine: ring of kCtaDoneRing per-launch "CTAs finished" counters
    uint32_t cta_done_next = 0;
    uint32_t* d_bitmat = nullptr;  // batched path: [queries][words] packed selection bits
    uint64_t bitmat_cap = 0;       // in words
    cudaStream_t side[2] = {nullptr, nullptr};  // batched path: phase-1 kernels alternate between two streams
    cudaEvent_t side_ev[3] = {nullptr, nullptr, nullptr};
    uint8_t* d_out = nullptr;
    uint64_t out_cap = 0;
    uint8_t* h_out = nullptr;  // pinned
    uint32_t* h_out_dev = nullptr;  // the same buffer as the device sees it (cudaHostGetDevicePointer)
    // One-launch single queries (AnswerPlan::final_out): every CTA XORs into this 4 KB accumulator and the launch's last
    // CTA moves the finished answer to its destination and leaves the accumulator zeroed.  It is all zeros between
    // launches; launches that use it are ordered by the stream (their global writes wait for the launch before,
    // answer_kernel.cu griddepcontrol.wait), and a call on another stream first waits for the last user (acc_ev).
    uint8_t* d_acc = nullptr;
    cudaStream_t acc_stream = nullptr;
    cudaEvent_t acc_ev = nullptr;
    // look-ahead (kernels.cuh LaMail / LaTop): key material of the launches in the queue, host-mapped; tops in HBM
    LaMail* h_mail = nullptr;
    const LaMail* h_mail_dev = nullptr;
    LaTop* d_tops = nullptr;
    LaMail* d_mail = nullptr;
    uint64_t la_seq = 0;
    // Host calls in flight (pir_answer from several threads): a ring of pinned, device-mapped answer slots, one event
    // each.  A caller enqueues its launch under the handle's mutex, then waits for ITS event with the mutex released,
    // so the launches of concurrent callers queue up behind each other on the GPU (and overlap / look ahead, 4.1).
    uint8_t* h_ring = nullptr;
    uint8_t* h_ring_dev = nullptr;
    cudaEvent_t ring_ev[kHostRing] = {};
    uint32_t ring_busy = 0;  // bit k: slot k belongs to a caller
    std::condition_variable ring_cv;
#ifdef PIRB_TEX_ROWS
    cudaTextureObject_t tex = 0;  // experiment: linear uint4 texture over d_slots (0 if the shard exceeds the texel limit)
#endif
    uint64_t h_out_cap = 0;
    // cached work split of the answer kernel for the last (num_bits, shard rows) seen
    uint32_t split_nb = 0, split_d = 0, split_hiw = 0, split_dmin = 0;
    uint64_t split_first = 0, split_rows = 0;
    bool split_tp = false;
    // optional per-launch timing of the dominant (fused) kernel, for bench.py's roofline
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pool;
    size_t ev_used = 0;
};

namespace {

int ensure_front(pir_db* db, uint64_t n) {
    if (n <= db->front_cap) return PIR_OK;
    for (int k = 0; k < 2; k++) {
        if (db->d_front_s[k]) cudaFree(db->d_front_s[k]);
        if (db->d_front_t[k]) cudaFree(db->d_front_t[k]);
        db->d_front_s[k] = nullptr;
        db->d_front_t[k] = nullptr;
    }
    db->front_cap = 0;
    for (int k = 0; k < 2; k++) {
        CU(cudaMalloc(&db->d_front_s[k], n * sizeof(uint4)));
        CU(cudaMalloc(&db->d_front_t[k], n));
    }
    db->front_cap = n;
    return PIR_OK;
}
int ensure_bits(pir_db* db, uint64_t n) {
    if (n <= db->bits_cap) return PIR_OK;
    if (db->d_bits) cudaFree(db->d_bits);
    db->d_bits = nullptr;
    db->bits_cap = 0;
    CU(cudaMalloc(&db->d_bits, n));
    db->bits_cap = n;
    return PIR_OK;
}
int ensure_out(pir_db* db, uint64_t n) {
    if (n > db->out_cap) {
        if (db->d_out) cudaFree(db->d_out);
        db->d_out = nullptr;
        db->out_cap = 0;
        CU(cudaMalloc(&db->d_out, n));
        db->out_cap = n;
    }
    if (n > db->h_out_cap) {
        if (db->h_out) cudaFreeHost(db->h_out);
        db->h_out = nullptr;
        db->h_out_cap = 0;
        db->h_out_dev = nullptr;
        CU(cudaMallocHost(&db->h_out, n));
        db->h_out_cap = n;
        void* dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, db->h_out, 0) == cudaSuccess) db->h_out_dev = (uint32_t*)dp;
        else cudaGetLastError();  // (no mapped access: the host calls keep the memset + copy path)
    }
    return PIR_OK;
}

// rows of the shard for group size W (global row indices)
int shard_rows(const pir_db* db, uint32_t W, uint64_t* first_row, uint64_t* n_rows) {
    if (W == 0) return fail(PIR_ERR_BAD_ARG, "group_size must be >= 1");
    uint64_t H = db->db_size / W;  // db.go:79,114: integer division
    if (db->slot_begin % W) return fail(PIR_ERR_BAD_ARG, "shard start %llu is not a multiple of group_size %u",
                                        (unsigned long long)db->slot_begin, W);
    uint64_t r0 = db->slot_begin / W;
    uint64_t end_slot = db->slot_begin + db->slot_count;
    if (end_slot != db->db_size && (db->slot_count % W))
        return fail(PIR_ERR_BAD_ARG, "shard length %llu is not a multiple of group_size %u",
                    (unsigned long long)db->slot_count, W);
    uint64_t r1 = end_slot / W;
    if (r1 > H) r1 = H;
    if (r0 > r1) r0 = r1;
    *first_row = r0;
    *n_rows = r1 - r0;
    return PIR_OK;
}

struct Plan {
    uint32_t D, B, Fc;
    uint64_t task_first, task_count;
};

// Choose the per-thread depth D and CTA breadth B (kernels.cuh) for a shard of rows.
Plan make_plan(uint32_t nb, uint64_t first_row, uint64_t n_rows, int num_sms, int forced = -1) {
    Plan best{};
    double best_score = -1.0;
    uint32_t Bmax = nb - 1 < (uint32_t)kFusedLogThreads ? nb - 1 : (uint32_t)kFusedLogThreads;
    for (int D = kMaxDfsDepth; D >= 0; D--) {
        if ((uint32_t)D + Bmax > nb - 1) continue;
        if (forced >= 0 && D != forced && (uint32_t)forced + Bmax <= nb - 1) continue;
        uint32_t B = Bmax;
        uint32_t Fc = nb - 1 - (uint32_t)D - B;
        uint32_t sh = nb - Fc;  // rows per task = 2^sh
        uint64_t t0 = first_row >> sh, t1 = (first_row + n_rows - 1) >> sh;
        uint64_t tasks = t1 - t0 + 1;
        uint64_t waves = (tasks + num_sms - 1) / num_sms;
        double balance = (double)tasks / (double)(waves * num_sms);
        double dfs = 16.0 * ((double)(2u << D) - 1.0);
        double score = balance * dfs / (dfs + 28.0);
        if (score > best_score) {
            best_score = score;
            best = Plan{(uint32_t)D, B, Fc, t0, tasks};
        }
    }
    return best;
}

// Work split of the answer kernel (answer_kernel.cu "work split"): uniform tasks of shape (D, hi_warps), or
// the binary split into aligned tasks of depth Dmin..D (hi_warps = 0).  Chosen against a cost model measured
// on B200 (tools/split_sweep.py, tools/live_warps_probe.py): a sub-partition of the SM expands one node for
// a warp in ~1.7k cycles, and with fewer than four busy warps it runs below that rate; the top of a uniform
// task costs ~1.3k cycles per tree level above the thread roots plus ~12k for its widest levels (34k at 17
// levels, 38k at 20); a task of the binary split costs ~36k on top of its nodes.
struct AnswerSplit {
    uint32_t D, hi_warps, Dmin;
};
double uniform_task_cost(uint32_t top_bits, uint32_t D, uint32_t hiw) {
    const double top = 1300.0 * (double)(top_bits - D) + 12000.0 + (hiw < 16u ? 2500.0 : 0.0);
    const double n_hi = (double)((2u << D) - 1u), n_lo = hiw < 16u ? (double)((1u << D) - 1u) : n_hi;
    const double pen = hiw >= 16u ? 1.0 : hiw == 12u ? 1.08 : hiw == 8u ? 1.25 : 1.7;
    return top + 1718.0 * (4.0 * n_lo + (double)(hiw / 4u) * (n_hi - n_lo) * pen);
}
// throughput: many independent launches run back to back on two streams (batches), so a CTA that finishes
// early is replaced by one of the next launch: total work matters more than the slowest CTA.
AnswerSplit choose_answer_split(uint32_t nb, uint64_t first_row, uint64_t n_rows, int num_sms, bool throughput = false,
                                const Tuning& tune = Tuning()) {
    AnswerSplit sp{};
    const uint32_t top_bits = nb - 1;
    uint32_t dmax = top_bits > (uint32_t)kFusedLogThreads ? top_bits - (uint32_t)kFusedLogThreads : 0u;
    if (dmax > 7u) dmax = 7u;
    if (tune.dfs_depth >= 0 && (uint32_t)tune.dfs_depth < dmax) dmax = (uint32_t)tune.dfs_depth;
    sp.D = dmax;
    sp.hi_warps = 16;
    // fixed splits (pir_db_set_option): dfs_depth x hi_warps (uniform), dfs_depth / dmin (binary)
    if (tune.hi_warps >= 0) {
        const uint32_t f = (uint32_t)tune.hi_warps;
        if (sp.D > 0 && (f == 4 || f == 8 || f == 12)) sp.hi_warps = f;
        return sp;
    }
    if (tune.dmin >= 0) {
        const uint32_t f = (uint32_t)tune.dmin;
        sp.hi_warps = 0;
        sp.Dmin = f < sp.D ? f : sp.D;
        return sp;
    }
    if (top_bits < (uint32_t)kFusedLogThreads + 1u) return sp;  // tiny domains: one uniform task of depth 0
    double best = -1.0;
    auto objective = [&](double worst, double sum) { return throughput ? 0.75 * sum / (double)num_sms + 0.25 * worst : worst; };
    for (uint32_t d = dmax + 1; d-- > 0;) {  // deep to shallow: a finer split must win by 0.5%
        for (uint32_t hiw = 16; hiw >= 4; hiw -= 4) {
            if (hiw < 16 && d == 0) break;
            const uint64_t nt = answer_tasks(first_row, n_rows, d, hiw);
            const double c = uniform_task_cost(top_bits, d, hiw);
            const double obj = objective(c * (double)((nt + (uint64_t)num_sms - 1) / (uint64_t)num_sms), c * (double)nt);
            if (best < 0.0 || obj < best * 0.995) {
                best = obj;
                sp.D = d;
                sp.hi_warps = hiw;
            }
        }
    }
    for (uint32_t dmin = dmax + 1; dmin-- > 0;) {  // binary split: replay every CTA's task sequence
        const uint64_t C = answer_chunks(first_row, n_rows, dmin);
        const uint64_t G = C < (uint64_t)num_sms ? C : (uint64_t)num_sms;
        const uint64_t cf = first_row >> (10 + dmin);
        double worst = 0.0, sum = 0.0;
        for (uint64_t b = 0; b < G; b++) {
            uint64_t cur = cf + b * C / G, end = cf + (b + 1) * C / G;
            double cost = 0.0;
            while (cur < end) {
                uint32_t d = dmin;
                uint64_t span = 1;
                while (d < dmax && !(cur & (2 * span - 1)) && cur + 2 * span <= end) {
                    d++;
                    span *= 2;
                }
                cost += 36000.0 + 4.0 * 1718.0 * (double)((2u << d) - 1u);
                cur += span;
            }
            if (cost > worst) worst = cost;
            sum += cost;
        }
        const double obj = objective(worst, sum);
        if (obj < best * 0.995) {
            best = obj;
            sp.D = dmax;
            sp.hi_warps = 0;
            sp.Dmin = dmin;
        }
    }
    return sp;
}
}  // namespace

extern "C" int pir_db_create(int device, uint64_t db_size, uint32_t slot_bytes, uint64_t slot_begin,
                             uint64_t slot_count, pir_db** out) {
    if (!out) return fail(PIR_ERR_BAD_ARG, "out is null");
    *out = nullptr;
    if (db_size == 0 || slot_bytes == 0) return fail(PIR_ERR_BAD_ARG, "db_size and slot_bytes must be non-zero");
    if (slot_begin > db_size || slot_count > db_size - slot_begin)
        return fail(PIR_ERR_BAD_ARG, "shard [%llu, +%llu) exceeds db_size %llu", (unsigned long long)slot_begin,
                    (unsigned long long)slot_count, (unsigned long long)db_size);
    DeviceCtx* ctx = nullptr;
    int rc = device_ctx(device, &ctx);
    if (rc) return rc;
    DeviceGuard g(device);
    pir_db* db = new (std::nothrow) pir_db();
    if (!db) return fail(PIR_ERR_OOM, "host allocation failed");
    db->device = device;
    db->ctx = ctx;
    db->db_size = db_size;
    db->slot_bytes = slot_bytes;
    db->pitch = (slot_bytes + 15u) & ~15u;
    db->slot_begin = slot_begin;
    db->slot_count = slot_count;
    uint64_t bytes = slot_count * (uint64_t)db->pitch;
    cudaError_t e = cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&db->d_slots, bytes ? bytes : 16);
    // Zero the slots (and the padding of slots that are not a multiple of 16 bytes).  On the handle's own stream and waited for:
    // cudaMemset() runs on the legacy default stream, may return before it has executed, and is NOT ordered against
    // db->stream (non-blocking) -- a fill or query enqueued right after pir_db_create could be overtaken by it and find
    // part of the database zeroed (seen as an intermittent wrong answer on 100-byte slots).
    // Every slot starts as zeros, like the reference's make([]byte, n) (slot.go:137-170): a query on a handle that has
    // been uploaded only in part must not XOR stale device memory into the answer.
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_slots, 0, bytes ? bytes : 16, db->stream);
    if (e == cudaSuccess) e = cudaMalloc(&db->d_acc, kAnswerMaxRowBytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_acc, 0, kAnswerMaxRowBytes, db->stream);
    if (e == cudaSuccess) e = cudaMalloc(&db->d_cta_done, kCtaDoneRing * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_cta_done, 0, kCtaDoneRing * sizeof(uint32_t), db->stream);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&db->acc_ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaHostAlloc(&db->h_ring, kHostRing * kAnswerMaxRowBytes, cudaHostAllocMapped);
    if (e == cudaSuccess) {
        void* dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, db->h_ring, 0) == cudaSuccess) db->h_ring_dev = (uint8_t*)dp;
        else cudaGetLastError();  // (no mapped access: host calls keep the memset + kernel + copy path)
    }
    for (uint32_t k = 0; k < kHostRing && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&db->ring_ev[k], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaHostAlloc(&db->h_mail, kLaMailRing * sizeof(LaMail), cudaHostAllocMapped);
    if (e == cudaSuccess) {
        memset(db->h_mail, 0, kLaMailRing * sizeof(LaMail));
        void* dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, db->h_mail, 0) == cudaSuccess) db->h_mail_dev = (const LaMail*)dp;
        else cudaGetLastError();  // (no mapped access: look-ahead stays off)
    }
    const size_t tops_bytes = (size_t)kLaTopRing * (size_t)ctx->num_sms * sizeof(LaTop);
    if (e == cudaSuccess) e = cudaMalloc(&db->d_tops, tops_bytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_tops, 0, tops_bytes, db->stream);
#ifdef PIRB_TEX_ROWS
    if (e == cudaSuccess && bytes) {
        cudaChannelFormatDesc cd = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindUnsigned);
        size_t max_texels = 0;
        if (cudaDeviceGetTexture1DLinearMaxWidth(&max_texels, &cd, device) != cudaSuccess) max_texels = 0, cudaGetLastError();
        if (bytes / 16 <= max_texels) {
            cudaResourceDesc rd{};
            rd.resType = cudaResourceTypeLinear;
            rd.res.linear.devPtr = db->d_slots;
            rd.res.linear.desc = cd;
            rd.res.linear.sizeInBytes = bytes;
            cudaTextureDesc td{};
            td.readMode = cudaReadModeElementType;
            td.filterMode = cudaFilterModePoint;
            td.addressMode[0] = cudaAddressModeClamp;
            e = cudaCreateTextureObject(&db->tex, &rd, &td, nullptr);
        }
        fprintf(stderr, "[tex] shard of %llu texels, limit %zu, texture %s\n", (unsigned long long)(bytes / 16), max_texels, db->tex ? "on" : "off");
    }
#endif
    if (e == cudaSuccess) e = cudaMalloc(&db->d_mail, kLaTopRing * sizeof(LaMail));
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_mail, 0, kLaTopRing * sizeof(LaMail), db->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(db->stream);
    if (e != cudaSuccess) {
        int code = e == cudaErrorMemoryAllocation ? PIR_ERR_OOM : PIR_ERR_CUDA;
        fail(code, "pir_db_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        pir_db_destroy(db);
        return code;
    }
    *out = db;
    return PIR_OK;
}

extern "C" int pir_db_destroy(pir_db* db) {
    if (!db) return PIR_OK;
    DeviceGuard g(db->device);
    if (db->stream) cudaStreamSynchronize(db->stream);
    for (int k = 0; k < 2; k++) {
        if (db->d_front_s[k]) cudaFree(db->d_front_s[k]);
        if (db->d_front_t[k]) cudaFree(db->d_front_t[k]);
    }
    if (db->d_bits) cudaFree(db->d_bits);
    if (db->d_bitmat) cudaFree(db->d_bitmat);
    if (db->d_cta_done) cudaFree(db->d_cta_done);
    if (db->d_acc) cudaFree(db->d_acc);
    if (db->acc_ev) cudaEventDestroy(db->acc_ev);
    if (db->h_ring) cudaFreeHost(db->h_ring);
    for (uint32_t k = 0; k < kHostRing; k++)
        if (db->ring_ev[k]) cudaEventDestroy(db->ring_ev[k]);
    if (db->d_tops) cudaFree(db->d_tops);
    if (db->d_mail) cudaFree(db->d_mail);
#ifdef PIRB_TEX_ROWS
    if (db->tex) cudaDestroyTextureObject(db->tex);
#endif
    if (db->h_mail) cudaFreeHost(db->h_mail);
    for (int k = 0; k < 2; k++)
        if (db->side[k]) cudaStreamDestroy(db->side[k]);
    for (int k = 0; k < 3; k++)
        if (db->side_ev[k]) cudaEventDestroy(db->side_ev[k]);
    if (db->d_out) cudaFree(db->d_out);
    if (db->h_out) cudaFreeHost(db->h_out);
    if (db->d_keywords) cudaFree(db->d_keywords);
    if (db->d_row_keywords) cudaFree(db->d_row_keywords);
    if (db->d_slots) cudaFree(db->d_slots);
    for (auto& ev : db->ev_pool) {
        cudaEventDestroy(ev.first);
        cudaEventDestroy(ev.second);
    }
    if (db->stream) cudaStreamDestroy(db->stream);
    delete db;
    return PIR_OK;
}

extern "C" int pir_db_set_kernel_timing(pir_db* db, int enable) {
    if (!db) return fail(PIR_ERR_BAD_ARG, "null handle");
    std::lock_guard<std::mutex> lk(db->mu);
    db->timing = enable != 0;
    db->ev_used = 0;
    return PIR_OK;
}

extern "C" int pir_db_get_kernel_timing(pir_db* db, double* total_ms, uint64_t* launches) {
    if (!db || !total_ms || !launches) return fail(PIR_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    double sum = 0.0;
    for (size_t i = 0; i < db->ev_used; i++) {
        CU(cudaEventSynchronize(db->ev_pool[i].second));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, db->ev_pool[i].first, db->ev_pool[i].second));
        sum += ms;
    }
    *total_ms = sum;
    *launches = db->ev_used;
    db->ev_used = 0;
    return PIR_OK;
}

namespace {
int set_tuning(Tuning& t, const char* name, int64_t v) {
    struct { const char* n; int* i; bool* b; int64_t lo, hi; } tab[] = {
        {"dfs_depth", &t.dfs_depth, nullptr, -1, 7},   {"hi_warps", &t.hi_warps, nullptr, -1, 16},
        {"dmin", &t.dmin, nullptr, -1, 7},             {"top_warp", &t.top_warp, nullptr, -1, 1},
        {"slice_log", &t.slice_log, nullptr, 0, 3},    {"batch_mode", &t.batch_mode, nullptr, 0, 2},
        {"batch_one_stream", nullptr, &t.batch_one_stream, 0, 1},
        {"batch_latency_split", nullptr, &t.batch_latency_split, 0, 1},
        {"kernel_v1", nullptr, &t.kernel_v1, 0, 1},    {"expand_v1", nullptr, &t.expand_v1, 0, 1},
        {"keyword_v1", nullptr, &t.keyword_v1, 0, 1},  {"force_unfused", nullptr, &t.force_unfused, 0, 1},
        {"signal_kernel", nullptr, &t.signal_kernel, 0, 1},
        {"host_copy_path", nullptr, &t.host_copy_path, 0, 1},
        {"serial_launch", nullptr, &t.serial_launch, 0, 1},
        {"no_lookahead", nullptr, &t.no_lookahead, 0, 1},
        {"no_trim", nullptr, &t.no_trim, 0, 1},
        {"direct_combine", nullptr, &t.direct_combine, 0, 1},
    };
    for (auto& e : tab) {
        if (strcmp(e.n, name)) continue;
        if (v < e.lo || v > e.hi) return fail(PIR_ERR_BAD_ARG, "option %s: value %lld outside [%lld, %lld]", name, (long long)v,
                                              (long long)e.lo, (long long)e.hi);
        if (e.i) *e.i = (int)v;
        else *e.b = v != 0;
        return PIR_OK;
    }
    return fail(PIR_ERR_BAD_ARG, "unknown option '%s'", name);
}
}  // namespace

extern "C" int pir_db_set_option(pir_db* db, const char* name, int64_t value) {
    if (!db || !name) return fail(PIR_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(db->mu);
    int rc = set_tuning(db->tune, name, value);
    db->split_nb = 0;  // the cached work split may no longer apply
    return rc;
}

extern "C" uint32_t pir_db_slot_pitch(const pir_db* db) { return db ? db->pitch : 0; }
extern "C" int pir_db_device(const pir_db* db) { return db ? db->device : -1; }

static int check_slot_range(const pir_db* db, uint64_t slot_index, uint64_t n_slots) {
    if (slot_index < db->slot_begin || slot_index - db->slot_begin > db->slot_count ||
        n_slots > db->slot_count - (slot_index - db->slot_begin))
        return fail(PIR_ERR_BAD_ARG, "slots [%llu, +%llu) are outside the shard", (unsigned long long)slot_index,
                    (unsigned long long)n_slots);
    return PIR_OK;
}

extern "C" int pir_db_upload(pir_db* db, uint64_t slot_index, const void* host_slots, uint64_t n_slots) {
    if (!db || (!host_slots && n_slots)) return fail(PIR_ERR_BAD_ARG, "null argument");
    int rc = check_slot_range(db, slot_index, n_slots);
    if (rc || n_slots == 0) return rc;
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    uint8_t* dst = db->d_slots + (slot_index - db->slot_begin) * (uint64_t)db->pitch;
    CU(cudaStreamSynchronize(db->stream));
    if (db->pitch == db->slot_bytes)
        CU(cudaMemcpy(dst, host_slots, n_slots * (uint64_t)db->slot_bytes, cudaMemcpyHostToDevice));
    else
        CU(cudaMemcpy2D(dst, db->pitch, host_slots, db->slot_bytes, db->slot_bytes, n_slots, cudaMemcpyHostToDevice));
    // (a copy from pageable memory may return once it is staged; the queries run on a non-blocking stream that the
    //  legacy default stream does not order, so wait for the copy itself)
    CU(cudaStreamSynchronize(0));
    return PIR_OK;
}

extern "C" int pir_db_download(pir_db* db, uint64_t slot_index, void* host_slots, uint64_t n_slots) {
    if (!db || (!host_slots && n_slots)) return fail(PIR_ERR_BAD_ARG, "null argument");
    int rc = check_slot_range(db, slot_index, n_slots);
    if (rc || n_slots == 0) return rc;
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    const uint8_t* src = db->d_slots + (slot_index - db->slot_begin) * (uint64_t)db->pitch;
    CU(cudaStreamSynchronize(db->stream));
    if (db->pitch == db->slot_bytes)
        CU(cudaMemcpy(host_slots, src, n_slots * (uint64_t)db->slot_bytes, cudaMemcpyDeviceToHost));
    else
        CU(cudaMemcpy2D(host_slots, db->slot_bytes, src, db->pitch, db->slot_bytes, n_slots, cudaMemcpyDeviceToHost));
    return PIR_OK;
}

extern "C" int pir_db_fill_synthetic(pir_db* db, uint64_t seed) {
    if (!db) return fail(PIR_ERR_BAD_ARG, "null handle");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    CU(launch_synth_fill(db->d_slots, db->slot_begin, db->slot_count, db->slot_bytes, db->pitch, seed, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    return PIR_OK;
}

extern "C" int pir_db_set_keywords(pir_db* db, const uint64_t* keywords, uint64_t n) {
    if (!db || !keywords) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (n != db->slot_count) return fail(PIR_ERR_BAD_ARG, "expected %llu keywords (one per shard slot), got %llu",
                                         (unsigned long long)db->slot_count, (unsigned long long)n);
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    if (!db->d_keywords) CU(cudaMalloc(&db->d_keywords, (n ? n : 1) * sizeof(uint64_t)));
    CU(cudaStreamSynchronize(db->stream));
    CU(cudaMemcpy(db->d_keywords, keywords, n * sizeof(uint64_t), cudaMemcpyHostToDevice));
    CU(cudaStreamSynchronize(0));  // (see pir_db_upload)
    return PIR_OK;
}

extern "C" int pir_db_set_row_keywords(pir_db* db, uint32_t group_size, const uint64_t* row_keywords, uint64_t n) {
    if (!db || (!row_keywords && n)) return fail(PIR_ERR_BAD_ARG, "null argument");
    uint64_t first = 0, rows = 0;
    int rc = shard_rows(db, group_size, &first, &rows);
    if (rc) return rc;
    if (n != rows) return fail(PIR_ERR_BAD_ARG, "expected %llu keywords (one per row of the shard at group size %u), got %llu",
                               (unsigned long long)rows, group_size, (unsigned long long)n);
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    CU(cudaStreamSynchronize(db->stream));
    if (db->d_row_keywords) cudaFree(db->d_row_keywords);
    db->d_row_keywords = nullptr;
    db->rowkw_w = 0;
    CU(cudaMalloc(&db->d_row_keywords, (n ? n : 1) * sizeof(uint64_t)));
    CU(cudaMemcpy(db->d_row_keywords, row_keywords, n * sizeof(uint64_t), cudaMemcpyHostToDevice));
    CU(cudaStreamSynchronize(0));  // (see pir_db_upload)
    db->rowkw_w = group_size;
    db->rowkw_first = first;
    db->rowkw_n = n;
    return PIR_OK;
}

extern "C" int pir_db_rows(const pir_db* db, uint32_t group_size, uint64_t* first_row, uint64_t* n_rows) {
    if (!db || !first_row || !n_rows) return fail(PIR_ERR_BAD_ARG, "null argument");
    return shard_rows(db, group_size, first_row, n_rows);
}

// ---- the answer path ----------------------------------------------------------------------------
namespace {
struct QueryGeom {
    uint32_t W;
    uint32_t nb;
    uint64_t first_row, n_rows;
    uint64_t row_bytes;  // pitched
};

void fill_answer_split(pir_db* db, const QueryGeom& g, AnswerPlan* ap, bool throughput = false);

// The keywords of the rows [g.first_row, g.first_row + g.n_rows) a keyword query evaluates on this handle (db.go:132-135:
// row r uses Keywords[r]).  With group size 1, and on an unsharded handle or a server's first shard, they are among the
// handle's per-slot keywords; a later shard queried with W > 1 needs the per-row copy of pir_db_set_row_keywords.
int shard_keywords(const pir_db* db, const QueryGeom& g, const uint64_t** kw) {
    if (!db->d_keywords) return fail(PIR_ERR_BAD_ARG, "keyword query on a database without keywords (pir_db_set_keywords)");
    if (g.first_row >= db->slot_begin && g.first_row + g.n_rows <= db->slot_begin + db->slot_count) {
        *kw = db->d_keywords + (g.first_row - db->slot_begin);
        return PIR_OK;
    }
    if (db->d_row_keywords && db->rowkw_w == g.W && db->rowkw_first == g.first_row && db->rowkw_n == g.n_rows) {
        *kw = db->d_row_keywords;
        return PIR_OK;
    }
    return fail(PIR_ERR_BAD_ARG,
                "keyword query with group_size %u on a shard that holds slots [%llu, %llu): its rows' keywords lie outside the shard; "
                "give them with pir_db_set_row_keywords (pir_server_set_keywords does, for the server's shard_align)",
                g.W, (unsigned long long)db->slot_begin, (unsigned long long)(db->slot_begin + db->slot_count));
}

int query_geom(const pir_db* db, const pir_query* q, QueryGeom* g) {
#ifdef PIRB_DEV
    {  // DEV build only: the tuning options follow the PIRB_* environment variables, re-read at every query
        pir_db* mdb = const_cast<pir_db*>(db);
        pirb_dev::dev_apply_env(mdb->tune);
        mdb->split_nb = 0;
    }
#endif
    if (!q) return fail(PIR_ERR_BAD_ARG, "query is null");
    if (q->reserved != 0) return fail(PIR_ERR_BAD_ARG, "pir_query.reserved must be 0");
    g->W = q->group_size;
    int rc = shard_rows(db, g->W, &g->first_row, &g->n_rows);
    if (rc) return rc;
    g->nb = pir_num_bits_for(db->db_size, g->W, q->is_keyword);
    if (q->num_bits != 0 && g->nb != 0 && q->num_bits != g->nb)
        return fail(PIR_ERR_BAD_KEY, "key has %u correction words but this database needs %u (db.go:114-121)",
                    q->num_bits, g->nb);
    g->row_bytes = (uint64_t)g->W * db->pitch;
    return PIR_OK;
}

// Enqueue one query's work on `st`; d_out receives the pitched answer (row_bytes bytes).

// arrivals / signalled: multi-GPU combine.  When the warp-specialised kernel answers the query, its last CTA
// bumps *arrivals itself (system scope, after its partial has been XORed into d_out) and *signalled is set;
// on every other path the caller launches the signalling kernel.
int enqueue_answer(pir_db* db, const pir_query* q, const QueryGeom& g, uint8_t* d_out, cudaStream_t st,
                   bool zero_first = true, uint32_t* arrivals = nullptr, bool* signalled = nullptr, bool throughput = false,
                   uint32_t* final_out = nullptr, const QueryConst* prebuilt = nullptr) {
    // every kernel below XOR-accumulates into d_out (atomics), so a caller that combines several shards
    // into one buffer passes zero_first = false and zeroes it once itself
    if (zero_first) CU(cudaMemsetAsync(d_out, 0, g.row_bytes, st));
    if (g.n_rows == 0) return PIR_OK;
    QueryConst qc_local;
    int rc = PIR_OK;
    if (!prebuilt && (rc = build_query_const(q, g.nb, &qc_local))) return rc;
    const QueryConst& qc = prebuilt ? *prebuilt : qc_local;  // (a multi-shard server marshals every key once, not once per shard)
    const DeviceCtx& ctx = *db->ctx;
    if (q->is_keyword) {
        const uint64_t* kw = nullptr;
        if ((rc = shard_keywords(db, g, &kw))) return rc;
        const uint8_t* rows0 = db->d_slots + (g.first_row * g.W - db->slot_begin) * (uint64_t)db->pitch;
        if (!db->tune.keyword_v1 && !db->tune.force_unfused && keyword_fused_supports(g.row_bytes)) {
            // one kernel: prefix-sharing walk with the row XOR fused in (no byte-per-row bit vector through HBM)
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (db->timing) {
                if (db->ev_used == db->ev_pool.size()) {
                    CU(cudaEventCreate(&e0));
                    CU(cudaEventCreate(&e1));
                    db->ev_pool.emplace_back(e0, e1);
                }
                e0 = db->ev_pool[db->ev_used].first;
                e1 = db->ev_pool[db->ev_used].second;
                db->ev_used++;
                CU(cudaEventRecord(e0, st));
            }
            CU(launch_keyword_walk(qc, kw, 0, g.n_rows, nullptr, nullptr, ctx.d_t0,
                                   ctx.smem_base, ctx.num_sms, st, rows0, g.row_bytes, (uint32_t*)d_out));
            if (e1) CU(cudaEventRecord(e1, st));
            return PIR_OK;
        }
        if ((rc = ensure_bits(db, g.n_rows))) return rc;
        if (db->tune.keyword_v1)
            CU(launch_eval_points(qc, kw, 0, g.n_rows, nullptr,
                                  db->d_bits, ctx.d_t0, ctx.smem_base, st));
        else
            CU(launch_keyword_walk(qc, kw, 0, g.n_rows, nullptr,
                                   db->d_bits, ctx.d_t0, ctx.smem_base, ctx.num_sms, st));
        CU(launch_scan_bytes(rows0, g.row_bytes, g.n_rows, db->d_bits, (uint32_t*)d_out, ctx.num_sms, st));
        return PIR_OK;
    }
    if (!db->tune.kernel_v1 && answer_kernel_supports(g.row_bytes, 0)) {
        // warp-specialised kernel: walks the top of the tree itself, one launch per query
        AnswerPlan ap{};
        ap.db = db->d_slots;
        ap.row_bytes = g.row_bytes;
        fill_answer_split(db, g, &ap, throughput);
#ifdef PIRB_TEX_ROWS
        ap.tex = db->tex;
        if (!db->tex) return fail(PIR_ERR_UNSUPPORTED, "PIRB_TEX_ROWS build: shard too large for a linear texture");
#endif
        ap.out = (uint32_t*)d_out;
        if (final_out) {  // (the caller has checked that this branch will be taken: single_launch_ok; d_out is db->d_acc)
            if (db->acc_stream && db->acc_stream != st) {
                // the accumulator's last user ran on another stream: this launch must not overlap it
                cudaError_t e = cudaEventRecord(db->acc_ev, db->acc_stream);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(st, db->acc_ev, 0);
                if (e != cudaSuccess) {  // (e.g. the caller has destroyed that stream since)
                    cudaGetLastError();
                    CU(cudaDeviceSynchronize());
                }
            }
            db->acc_stream = st;
            ap.final_out = final_out;
            ap.cta_done = db->d_cta_done + (db->cta_done_next++ % kCtaDoneRing);
            if (arrivals && signalled) {  // multi-GPU: the last CTA XORs the shard's sum into final_out (a combine slot) and signals
                ap.arrivals = arrivals;
                *signalled = true;
            }
        } else if (arrivals && signalled && !db->tune.signal_kernel) {
            ap.arrivals = arrivals;
            ap.cta_done = db->d_cta_done + (db->cta_done_next++ % kCtaDoneRing);  // reset by the launch's last CTA
            *signalled = true;
        }
        // Look-ahead: leave this launch's key material where the launch before it can find it, and tell this launch where
        // to look for the one after it.  Only for launches of one task per CTA with a walker warp (rows >= 128 B).
        if (!db->tune.no_lookahead && db->h_mail_dev && ap.top_warp && ap.hi_warps != 0u && g.nb - 1u - ap.D <= kLaMaxLevels &&
            answer_tasks(ap.first_row, ap.n_rows, ap.D, ap.hi_warps, ap.m_task) <= (uint64_t)ctx.num_sms) {
            const uint64_t seq = ++db->la_seq;
            uint64_t hsh = 0x9E3779B97F4A7C15ull;
            for (uint64_t v : {(uint64_t)g.nb, (uint64_t)ap.D, (uint64_t)ap.hi_warps, (uint64_t)ap.m_task, (uint64_t)ap.slog_cap, ap.first_row,
                               ap.n_rows, (uint64_t)g.row_bytes})
                hsh = (hsh ^ v) * 0xBF58476D1CE4E5B9ull + (hsh >> 29);
            LaMail* m = db->h_mail + (seq % kLaMailRing);
            volatile uint64_t* tag = &m->tag;
            *tag = 0;  // seqlock: invalid while the payload changes (a launch up to 64 back may still poll this entry)
            std::atomic_thread_fence(std::memory_order_seq_cst);
            m->geom = hsh;
            memcpy(m->c.rk, qc.rk, sizeof m->c.rk);
            memcpy(m->c.cw, qc.cw, sizeof m->c.cw);
            m->c.cw_tl = qc.cw_tl;
            m->c.cw_tr = qc.cw_tr;
            memcpy(m->c.s_init, qc.s_init, sizeof m->c.s_init);
            m->c.t_init = qc.t_init;
            std::atomic_thread_fence(std::memory_order_seq_cst);
            *tag = seq;
            ap.la_seq = seq;
            ap.la_geom = hsh;
            ap.la_mail = db->h_mail_dev;
            ap.la_dmail = db->d_mail;
            ap.la_tops = db->d_tops;
        }
#ifdef PIRB_DEV
        DevPhase dev_phase;
        if (int drc = dev_phase.begin(&ap, st)) return drc;
#endif
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (db->timing) {
            if (db->ev_used == db->ev_pool.size()) {
                CU(cudaEventCreate(&e0));
                CU(cudaEventCreate(&e1));
                db->ev_pool.emplace_back(e0, e1);
            }
            e0 = db->ev_pool[db->ev_used].first;
            e1 = db->ev_pool[db->ev_used].second;
            db->ev_used++;
            CU(cudaEventRecord(e0, st));
        }
        // consecutive answer launches overlap (programmatic dependent launch) unless a launch is being timed on its own
        CU(launch_answer(qc, ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, st, /*overlap=*/!db->timing && !db->tune.serial_launch));
        if (e1) CU(cudaEventRecord(e1, st));
#ifdef PIRB_DEV
        if (int drc = dev_phase.report(st)) return drc;
#endif
        return PIR_OK;
    }
    Plan p = make_plan(g.nb, g.first_row, g.n_rows, ctx.num_sms, db->tune.dfs_depth);
    if ((rc = ensure_front(db, p.task_count + 2))) return rc;
    TopPlan tp{};
    tp.Fc = p.Fc;
    tp.task_first = p.task_first;
    tp.task_count = p.task_count;
    for (int k = 0; k < 2; k++) {
        tp.buf_s[k] = db->d_front_s[k];
        tp.buf_t[k] = db->d_front_t[k];
    }
    CU(launch_top(qc, tp, ctx.d_t0, ctx.smem_base, st));
    FusedPlan fp{};
    fp.db = db->d_slots;
    fp.row_bytes = g.row_bytes;
    fp.first_row = g.first_row;
    fp.n_rows = g.n_rows;
    fp.D = p.D;
    fp.B = p.B;
    fp.Fc = p.Fc;
    fp.task_first = p.task_first;
    fp.task_count = p.task_count;
    fp.frontier_s = db->d_front_s[p.Fc & 1];
    fp.frontier_t = db->d_front_t[p.Fc & 1];
    fp.out = (uint32_t*)d_out;
    const bool fused = fused_supports_row_bytes(g.row_bytes) && !db->tune.force_unfused;
    if (fused) {
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (db->timing) {
            if (db->ev_used == db->ev_pool.size()) {
                CU(cudaEventCreate(&e0));
                CU(cudaEventCreate(&e1));
                db->ev_pool.emplace_back(e0, e1);
            }
            e0 = db->ev_pool[db->ev_used].first;
            e1 = db->ev_pool[db->ev_used].second;
            db->ev_used++;
            CU(cudaEventRecord(e0, st));
        }
        CU(launch_fused(qc, fp, false, ctx.d_t0, ctx.smem_base, ctx.num_sms, st));
        if (e1) CU(cudaEventRecord(e1, st));
    } else {
        if ((rc = ensure_bits(db, g.n_rows))) return rc;
        fp.bits_out = db->d_bits;
        fp.row_bytes = 16;  // unused in BITS mode; keeps the kernel's row geometry trivial
        CU(launch_fused(qc, fp, true, ctx.d_t0, ctx.smem_base, ctx.num_sms, st));
        CU(launch_scan_bytes(db->d_slots, g.row_bytes, g.n_rows, db->d_bits, (uint32_t*)d_out, ctx.num_sms, st));
    }
    return PIR_OK;
}

// Fill the work-split fields of an AnswerPlan (cached per handle for the last (num_bits, rows) seen).
void fill_answer_split(pir_db* db, const QueryGeom& g, AnswerPlan* ap, bool throughput) {
    if (db->split_nb != g.nb || db->split_first != g.first_row || db->split_rows != g.n_rows || db->split_tp != throughput) {
        const AnswerSplit sp = choose_answer_split(g.nb, g.first_row, g.n_rows, db->ctx->num_sms, throughput, db->tune);
        db->split_tp = throughput;
        db->split_nb = g.nb;
        db->split_first = g.first_row;
        db->split_rows = g.n_rows;
        db->split_d = sp.D;
        db->split_hiw = sp.hi_warps;
        db->split_dmin = sp.Dmin;
    }
    ap->first_row = g.first_row;
    ap->n_rows = g.n_rows;
    ap->D = db->split_d;
    ap->hi_warps = db->split_hiw;
    ap->Dmin = db->split_dmin;
    ap->m_task = 0;
    if (ap->hi_warps != 0u && !throughput && !db->tune.no_trim) {
        // Trim the tasks so that every SM gets the same number of rows: 2^24 rows in (7, 12) tasks of 114 688 rows are
        // 146.3 tasks, i.e. 147 CTAs of which one is a third full and one SM idle; 148 tasks of 443 (instead of 448)
        // thread-root nodes do 1 % less work each.  Only a light trim: the trailing threads of a task idle.
        const uint64_t t0 = g.first_row & ~((2ull << ap->D) - 1ull);
        const uint64_t nodes = (g.first_row + g.n_rows - t0 + (2ull << ap->D) - 1ull) >> (ap->D + 1u);
        const uint64_t nom = answer_task_nodes(ap->hi_warps, 0);
        const uint64_t nt = (nodes + nom - 1) / nom, sms = (uint64_t)db->ctx->num_sms;
        const uint64_t slots = (nt + sms - 1) / sms * sms;
        const uint64_t m = (nodes + slots - 1) / slots;
        if (m < nom && 8 * m >= 7 * nom) ap->m_task = (uint32_t)m;
    }
    // Slices of 8 rows per thread, or of 4 when the task a CTA ends with is shallow: the rows of the last slice
    // are scanned after the tree warps are done, and with 32 rows per thread that slice is a quarter of the task.
    ap->slog_cap = (ap->hi_warps != 0u && ap->D <= 4u) ? 2u : 3u;
    // The last scan warp walks the tops of the tasks ahead (answer_kernel.cu TopWalker).  It pays when the scan warps
    // have slack: 2^24 x 256 B in several tasks per CTA 0.943 -> 0.937 ms; with 32-byte rows they have none
    // (2^28 x 32 B: 13.05 ms without the walker, 13.22 ms with it).
    ap->top_warp = g.row_bytes >= 128u ? 1u : 0u;
    if (db->tune.top_warp >= 0) ap->top_warp = db->tune.top_warp ? 1u : 0u;
    if (db->tune.slice_log == 2 || db->tune.slice_log == 3) ap->slog_cap = (uint32_t)db->tune.slice_log;
}

// Should a batch of n index queries go through the two-phase path (bit vectors, then ONE pass over the
// rows for all queries)?  It pays when the fused kernel would be bound by streaming the rows once per
// query, i.e. for wide rows; PIRB_BATCH_MODE=fused|bits overrides.
bool batch_uses_bits_path(const pir_db* db, const pir_query* qs, uint32_t n, const QueryGeom& g) {
    if (qs[0].is_keyword || g.n_rows == 0) return false;
    uint32_t tpq, segc, sr, qpl;
    if (!batch_scan_geometry(g.row_bytes, &tpq, &segc, &sr, &qpl)) return false;
    if (db->tune.batch_mode == 1) return false;
    if (db->tune.batch_mode == 2) return true;
    return n >= 4 && g.row_bytes >= 512;
}

// Two-phase batch: phase 1 expands every query into a packed bit vector (answer_kernel, BITS mode: no row
// is read), phase 2 streams the shard once and folds every row into the accumulators of the queries that
// selected it (batch_scan.cu).  d_out: n pitched answers, XOR-accumulated (zeroed here if zero_first).
int enqueue_batch_bits(pir_db* db, const pir_query* qs, uint32_t n, const QueryGeom& g, uint8_t* d_out, cudaStream_t st,
                       bool zero_first = true, const QueryConst* prebuilt = nullptr) {
    // Every key is validated and marshalled BEFORE anything is enqueued: a bad key at position k must not leave the
    // kernels of queries 0..k-1 running on the side streams behind the caller's back.
    std::vector<QueryConst> qcs_local;
    if (!prebuilt) {
        qcs_local.resize(n);
        for (uint32_t i = 0; i < n; i++) {
            int rc = build_query_const(&qs[i], g.nb, &qcs_local[i]);
            if (rc) return rc;
        }
    }
    const QueryConst* qcs = prebuilt ? prebuilt : qcs_local.data();
    const DeviceCtx& ctx = *db->ctx;
    const uint64_t wpq = (((g.n_rows + 31) >> 5) + 1 + 3) & ~3ull;  // +1: a slice may spill into the next word
    if ((uint64_t)n * wpq > db->bitmat_cap) {
        if (db->d_bitmat) cudaFree(db->d_bitmat);
        db->d_bitmat = nullptr;
        db->bitmat_cap = 0;
        CU(cudaMalloc(&db->d_bitmat, (uint64_t)n * wpq * 4));
        db->bitmat_cap = (uint64_t)n * wpq;
    }
    CU(cudaMemsetAsync(db->d_bitmat, 0, (uint64_t)n * wpq * 4, st));
    if (zero_first) CU(cudaMemsetAsync(d_out, 0, (uint64_t)n * g.row_bytes, st));
    // Phase-1 kernels are independent; alternating them between two streams lets the CTAs of the next
    // kernel start on SMs that the previous one has already left (each CTA fills an SM, so they never
    // share one), which hides the ramp-down of every launch.
    const bool two_streams = n >= 2 && !db->tune.batch_one_stream;
    if (two_streams) {
        for (int k = 0; k < 2; k++)
            if (!db->side[k]) CU(cudaStreamCreateWithFlags(&db->side[k], cudaStreamNonBlocking));
        for (int k = 0; k < 3; k++)
            if (!db->side_ev[k]) CU(cudaEventCreateWithFlags(&db->side_ev[k], cudaEventDisableTiming));
        CU(cudaEventRecord(db->side_ev[2], st));
        for (int k = 0; k < 2; k++) CU(cudaStreamWaitEvent(db->side[k], db->side_ev[2], 0));
    }
    cudaError_t lerr = cudaSuccess;
    for (uint32_t i = 0; i < n && lerr == cudaSuccess; i++) {
        AnswerPlan ap{};
        ap.db = db->d_slots;
        ap.row_bytes = 16;  // unused in BITS mode
        fill_answer_split(db, g, &ap, /*throughput=*/two_streams && n >= 4);
        ap.bits_out = db->d_bitmat + (uint64_t)i * wpq;
        lerr = launch_answer(qcs[i], ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, two_streams ? db->side[i & 1] : st);
    }
    if (two_streams)
        for (int k = 0; k < 2; k++) {  // join even on failure, so that `st` never runs ahead of the side streams
            cudaError_t e1 = cudaEventRecord(db->side_ev[k], db->side[k]);
            if (e1 == cudaSuccess) e1 = cudaStreamWaitEvent(st, db->side_ev[k], 0);
            if (lerr == cudaSuccess) lerr = e1;
        }
    CU(lerr);
    uint32_t tpq, segc, sr, qpl;
    batch_scan_geometry(g.row_bytes, &tpq, &segc, &sr, &qpl);
#ifdef PIRB_DEV
    if (pirb_dev::dev_skip_scan()) return PIR_OK;  // timing experiments only (tools/batch_phases.py)
#endif
    for (uint32_t q0 = 0; q0 < n; q0 += qpl) {
        BatchScanPlan bp{};
        bp.db = db->d_slots;
        bp.row_bytes = g.row_bytes;
        bp.n_rows = g.n_rows;
        bp.bits = db->d_bitmat + (uint64_t)q0 * wpq;
        bp.wpq = wpq;
        bp.n_queries = n - q0 < qpl ? n - q0 : qpl;
        bp.tpq = tpq;
        bp.seg_chunks = segc;
        bp.stage_rows = sr;
        bp.rsplit = 1;
        while (bp.rsplit < 32 && (uint64_t)bp.n_queries * tpq * (bp.rsplit * 2) <= 512) bp.rsplit *= 2;
        bp.out = (uint32_t*)(d_out + (uint64_t)q0 * g.row_bytes);
        CU(launch_batch_scan(bp, ctx.num_sms, st));
    }
    return PIR_OK;
}

// pitched device answer -> dense host answer (Shares[c].Data = out[c*S, (c+1)*S))
// A batch answered by one fused launch per query (rows narrower than 512 B).  The launches are independent (each
// XORs into its own answer), so they alternate between two side streams: the CTAs of the next launch start on SMs
// that the previous one has already left, which hides every launch's ramp-down and launch gap.  Keyword queries
// and the two-launch fallback share per-handle scratch and stay on one stream.
int enqueue_batch_fused(pir_db* db, const pir_query* qs, uint32_t n, const QueryGeom& g0, uint8_t* d_out, cudaStream_t st) {
    int rc;
    const bool two_streams = n >= 2 && !qs[0].is_keyword && !db->tune.kernel_v1 && g0.n_rows != 0 &&
                             answer_kernel_supports(g0.row_bytes, 0) && !db->tune.batch_one_stream && !dev_phase_debug_on();
    if (!two_streams) {
        for (uint32_t i = 0; i < n; i++)
            if ((rc = enqueue_answer(db, &qs[i], g0, d_out + (uint64_t)i * g0.row_bytes, st))) return rc;
        return PIR_OK;
    }
    for (int k = 0; k < 2; k++)
        if (!db->side[k]) CU(cudaStreamCreateWithFlags(&db->side[k], cudaStreamNonBlocking));
    for (int k = 0; k < 3; k++)
        if (!db->side_ev[k]) CU(cudaEventCreateWithFlags(&db->side_ev[k], cudaEventDisableTiming));
    CU(cudaMemsetAsync(d_out, 0, (uint64_t)n * g0.row_bytes, st));
    CU(cudaEventRecord(db->side_ev[2], st));
    for (int k = 0; k < 2; k++) CU(cudaStreamWaitEvent(db->side[k], db->side_ev[2], 0));
    rc = PIR_OK;
    for (uint32_t i = 0; i < n && rc == PIR_OK; i++)
        rc = enqueue_answer(db, &qs[i], g0, d_out + (uint64_t)i * g0.row_bytes, db->side[i & 1], /*zero_first=*/false, nullptr,
                            nullptr, /*throughput=*/n >= 4 && !db->tune.batch_latency_split);
    for (int k = 0; k < 2; k++) {  // join even on failure, so that `st` never runs ahead of the side streams
        CU(cudaEventRecord(db->side_ev[k], db->side[k]));
        CU(cudaStreamWaitEvent(st, db->side_ev[k], 0));
    }
    return rc;
}

// Can a single query be answered by ONE launch of the warp-specialised kernel (AnswerPlan::final_out)?
bool single_launch_ok(const pir_db* db, const pir_query* q, const QueryGeom& g) {
    return !db->tune.host_copy_path && !db->tune.kernel_v1 && !q->is_keyword && g.n_rows != 0 &&
           answer_kernel_supports(g.row_bytes, 0) && g.row_bytes <= kAnswerMaxRowBytes && !dev_phase_debug_on();
}

void compact_answer(const pir_db* db, uint32_t W, const uint8_t* pitched, uint8_t* dense) {
    for (uint32_t c = 0; c < W; c++) memcpy(dense + (size_t)c * db->slot_bytes, pitched + (size_t)c * db->pitch, db->slot_bytes);
}
}  // namespace

extern "C" int pir_answer_batch_device(pir_db* db, const pir_query* qs, uint32_t n_queries, void* d_out,
                                       void* stream) {
    if (!db || !qs || !d_out || n_queries == 0) return fail(PIR_ERR_BAD_ARG, "null or empty argument");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    cudaStream_t st = stream ? (cudaStream_t)stream : db->stream;
    QueryGeom g0;
    int rc = query_geom(db, &qs[0], &g0);
    if (rc) return rc;
    for (uint32_t i = 0; i < n_queries; i++) {
        QueryGeom geom;
        if ((rc = query_geom(db, &qs[i], &geom))) return rc;
        if (qs[i].group_size != qs[0].group_size || qs[i].is_keyword != qs[0].is_keyword)
            return fail(PIR_ERR_BAD_ARG, "all queries of a batch must share group_size and is_keyword");
    }
    // One index query: one launch, no memset (the last CTA moves the answer from the handle's accumulator to d_out),
    // so back-to-back calls on one stream overlap as programmatic dependent launches.
    if (n_queries == 1 && single_launch_ok(db, &qs[0], g0))
        return enqueue_answer(db, &qs[0], g0, db->d_acc, st, /*zero_first=*/false, nullptr, nullptr, false, (uint32_t*)d_out);
    if (batch_uses_bits_path(db, qs, n_queries, g0)) return enqueue_batch_bits(db, qs, n_queries, g0, (uint8_t*)d_out, st);
    return enqueue_batch_fused(db, qs, n_queries, g0, (uint8_t*)d_out, st);
}

extern "C" int pir_answer_batch(pir_db* db, const pir_query* qs, uint32_t n_queries, void* out, size_t out_len) {
    if (!db || !qs || !out || n_queries == 0) return fail(PIR_ERR_BAD_ARG, "null or empty argument");
    std::unique_lock<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    QueryGeom g0;
    int rc = query_geom(db, &qs[0], &g0);
    if (rc) return rc;
    const size_t ans = (size_t)g0.W * db->slot_bytes;
    if (out_len < ans * n_queries) return fail(PIR_ERR_BAD_ARG, "out_len %zu < %zu", out_len, ans * n_queries);
    // One index query answered by the warp-specialised kernel: its last CTA writes the answer straight into a pinned host
    // slot and re-zeroes the device accumulator, so the call is one launch and one wait -- for the caller's OWN event,
    // with the handle's mutex released: calls from several threads are in flight together (kHostRing at most).
    if (n_queries == 1 && db->h_ring_dev && single_launch_ok(db, &qs[0], g0)) {
        db->ring_cv.wait(lk, [&] { return db->ring_busy != (1u << kHostRing) - 1u; });
        uint32_t k = 0;
        while ((db->ring_busy >> k) & 1u) k++;
        db->ring_busy |= 1u << k;
        uint8_t* slot = db->h_ring + (size_t)k * kAnswerMaxRowBytes;
        rc = enqueue_answer(db, &qs[0], g0, db->d_acc, db->stream, /*zero_first=*/false, nullptr, nullptr, false,
                            (uint32_t*)(db->h_ring_dev + (size_t)k * kAnswerMaxRowBytes));
        cudaError_t e = cudaEventRecord(db->ring_ev[k], db->stream);
        lk.unlock();
        if (e == cudaSuccess) e = cudaEventSynchronize(db->ring_ev[k]);
        if (e != cudaSuccess && rc == PIR_OK) rc = fail(PIR_ERR_CUDA, "answer kernel failed: %s", cudaGetErrorString(e));
        if (rc == PIR_OK) compact_answer(db, g0.W, slot, (uint8_t*)out);
        lk.lock();
        db->ring_busy &= ~(1u << k);
        db->ring_cv.notify_one();
        return rc;
    }
    if ((rc = ensure_out(db, g0.row_bytes * n_queries))) return rc;
    for (uint32_t i = 0; i < n_queries; i++) {
        QueryGeom geom;
        if ((rc = query_geom(db, &qs[i], &geom))) return rc;
        if (qs[i].group_size != qs[0].group_size || qs[i].is_keyword != qs[0].is_keyword)
            return fail(PIR_ERR_BAD_ARG, "all queries of a batch must share group_size and is_keyword");
    }
    if (batch_uses_bits_path(db, qs, n_queries, g0)) {
        if ((rc = enqueue_batch_bits(db, qs, n_queries, g0, db->d_out, db->stream))) {
            cudaStreamSynchronize(db->stream);
            return rc;
        }
    } else {
        if ((rc = enqueue_batch_fused(db, qs, n_queries, g0, db->d_out, db->stream))) {
            cudaStreamSynchronize(db->stream);
            return rc;
        }
    }
    CU(cudaMemcpyAsync(db->h_out, db->d_out, g0.row_bytes * n_queries, cudaMemcpyDeviceToHost, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    for (uint32_t i = 0; i < n_queries; i++)
        compact_answer(db, g0.W, db->h_out + (uint64_t)i * g0.row_bytes, (uint8_t*)out + (size_t)i * ans);
    return PIR_OK;
}

extern "C" int pir_answer(pir_db* db, const pir_query* q, void* out, size_t out_len) {
    return pir_answer_batch(db, q, 1, out, out_len);
}

extern "C" int pir_expand_bits(pir_db* db, const pir_query* q, uint8_t* bits_out, uint64_t n_rows) {
    if (!db || !q || (!bits_out && n_rows)) return fail(PIR_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    QueryGeom geom;
    int rc = query_geom(db, q, &geom);
    if (rc) return rc;
    if (n_rows != geom.n_rows) return fail(PIR_ERR_BAD_ARG, "n_rows %llu != shard rows %llu",
                                           (unsigned long long)n_rows, (unsigned long long)geom.n_rows);
    if (n_rows == 0) return PIR_OK;
    QueryConst qc;
    if ((rc = build_query_const(q, geom.nb, &qc))) return rc;
    if ((rc = ensure_bits(db, n_rows))) return rc;
    const DeviceCtx& ctx = *db->ctx;
    if (q->is_keyword) {
        const uint64_t* kw = nullptr;
        if ((rc = shard_keywords(db, geom, &kw))) return rc;
        if (db->tune.keyword_v1)
            CU(launch_eval_points(qc, kw, 0, n_rows, nullptr,
                                  db->d_bits, ctx.d_t0, ctx.smem_base, db->stream));
        else
            CU(launch_keyword_walk(qc, kw, 0, n_rows, nullptr,
                                   db->d_bits, ctx.d_t0, ctx.smem_base, ctx.num_sms, db->stream));
    } else if (!db->tune.expand_v1) {
        // answer kernel in BITS mode (packed bit vector, no rows read), then one byte per row
        const uint64_t wpq = (((n_rows + 31) >> 5) + 1 + 3) & ~3ull;
        if (wpq > db->bitmat_cap) {
            if (db->d_bitmat) cudaFree(db->d_bitmat);
            db->d_bitmat = nullptr;
            db->bitmat_cap = 0;
            CU(cudaMalloc(&db->d_bitmat, wpq * 4));
            db->bitmat_cap = wpq;
        }
        CU(cudaMemsetAsync(db->d_bitmat, 0, wpq * 4, db->stream));
        AnswerPlan ap{};
        ap.db = db->d_slots;
        ap.row_bytes = 16;
        fill_answer_split(db, geom, &ap);
        ap.bits_out = db->d_bitmat;
        CU(launch_answer(qc, ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, db->stream));
        CU(launch_unpack_bits(db->d_bitmat, db->d_bits, n_rows, db->stream));
    } else {
        Plan p = make_plan(geom.nb, geom.first_row, geom.n_rows, ctx.num_sms, db->tune.dfs_depth);
        if ((rc = ensure_front(db, p.task_count + 2))) return rc;
        TopPlan tp{};
        tp.Fc = p.Fc;
        tp.task_first = p.task_first;
        tp.task_count = p.task_count;
        for (int k = 0; k < 2; k++) {
            tp.buf_s[k] = db->d_front_s[k];
            tp.buf_t[k] = db->d_front_t[k];
        }
        CU(launch_top(qc, tp, ctx.d_t0, ctx.smem_base, db->stream));
        FusedPlan fp{};
        fp.db = db->d_slots;
        fp.row_bytes = 16;
        fp.first_row = geom.first_row;
        fp.n_rows = geom.n_rows;
        fp.D = p.D;
        fp.B = p.B;
        fp.Fc = p.Fc;
        fp.task_first = p.task_first;
        fp.task_count = p.task_count;
        fp.frontier_s = db->d_front_s[p.Fc & 1];
        fp.frontier_t = db->d_front_t[p.Fc & 1];
        fp.bits_out = db->d_bits;
        CU(launch_fused(qc, fp, true, ctx.d_t0, ctx.smem_base, ctx.num_sms, db->stream));
    }
    CU(cudaMemcpyAsync(bits_out, db->d_bits, n_rows, cudaMemcpyDeviceToHost, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    return PIR_OK;
}

// The AES-only rate of this shard, measured: `reps` launches of the answer kernel in BITS mode (the same tree
// expansion for the same work split, selection bits OR-ed into a packed vector, no database row read) between two
// CUDA events on the handle's stream.  bench.py's lookup-pipe roofline divides by this.
extern "C" int pir_db_time_expansion(pir_db* db, const pir_query* q, uint32_t reps, double* ms_per_launch) {
    if (!db || !q || !ms_per_launch || reps == 0) return fail(PIR_ERR_BAD_ARG, "null or empty argument");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    QueryGeom geom;
    int rc = query_geom(db, q, &geom);
    if (rc) return rc;
    if (q->is_keyword || geom.n_rows == 0) return fail(PIR_ERR_BAD_ARG, "needs an index query over a non-empty shard");
    QueryConst qc;
    if ((rc = build_query_const(q, geom.nb, &qc))) return rc;
    const DeviceCtx& ctx = *db->ctx;
    const uint64_t wpq = (((geom.n_rows + 31) >> 5) + 1 + 3) & ~3ull;
    if (wpq > db->bitmat_cap) {
        if (db->d_bitmat) cudaFree(db->d_bitmat);
        db->d_bitmat = nullptr;
        db->bitmat_cap = 0;
        CU(cudaMalloc(&db->d_bitmat, wpq * 4));
        db->bitmat_cap = wpq;
    }
    AnswerPlan ap{};
    ap.db = db->d_slots;
    ap.row_bytes = 16;
    fill_answer_split(db, geom, &ap);
    ap.bits_out = db->d_bitmat;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    CU(cudaEventCreate(&e0));
    cudaError_t e = cudaEventCreate(&e1);
    if (e == cudaSuccess) e = cudaMemsetAsync(db->d_bitmat, 0, wpq * 4, db->stream);
    if (e == cudaSuccess) e = launch_answer(qc, ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, db->stream);  // warm-up
    if (e == cudaSuccess) e = cudaEventRecord(e0, db->stream);
    for (uint32_t i = 0; i < reps && e == cudaSuccess; i++) e = launch_answer(qc, ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, db->stream);
    if (e == cudaSuccess) e = cudaEventRecord(e1, db->stream);
    if (e == cudaSuccess) e = cudaEventSynchronize(e1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (e != cudaSuccess) {
        cudaStreamSynchronize(db->stream);
        return fail(PIR_ERR_CUDA, "pir_db_time_expansion: %s", cudaGetErrorString(e));
    }
    *ms_per_launch = (double)ms / (double)reps;
    return PIR_OK;
}

// ExpandSharedQuery without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The same two kernels as pir_expand_bits -- answer_kernel in BITS mode, then unpack_bits_kernel -- with the packed bit
// vector in the static device array g_noalloc_bits instead of per-handle scratch from cudaMalloc, and the result
// written to the CALLER's device buffer.  Only enqueues work on `stream`: no cudaMalloc / cudaFree / event or stream
// creation, no synchronisation.  Every call on a GPU shares the one array: it is cleared at the start of every call
// (the kernel ORs into it), and two calls that overlap on the GPU race.
extern "C" int pir_expand_bits_noalloc(pir_db* db, const pir_query* q, uint8_t* d_bits_out, uint64_t n_rows, void* stream) {
    if (!db || !q || (!d_bits_out && n_rows)) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (q->is_keyword) return fail(PIR_ERR_UNSUPPORTED, "pir_expand_bits_noalloc takes index queries only (keyword queries use no scratch)");
    if (((uintptr_t)d_bits_out & 3u) != 0) return fail(PIR_ERR_BAD_ARG, "d_bits_out must be 4-byte aligned");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    QueryGeom geom;
    int rc = query_geom(db, q, &geom);
    if (rc) return rc;
    if (n_rows != geom.n_rows) return fail(PIR_ERR_BAD_ARG, "n_rows %llu != shard rows %llu", (unsigned long long)n_rows,
                                           (unsigned long long)geom.n_rows);
    if (n_rows > kNoAllocMaxRows)
        return fail(PIR_ERR_BAD_ARG, "pir_expand_bits_noalloc: %llu rows exceed the static scratch (PIR_NOALLOC_MAX_ROWS = %llu); "
                                     "use pir_expand_bits", (unsigned long long)n_rows, (unsigned long long)kNoAllocMaxRows);
    if (n_rows == 0) return PIR_OK;
    QueryConst qc;
    if ((rc = build_query_const(q, geom.nb, &qc))) return rc;
    const DeviceCtx& ctx = *db->ctx;
    if (!ctx.noalloc_bits) return fail(PIR_ERR_CUDA, "static scratch was not set up on device %d", db->device);
    cudaStream_t st = stream ? (cudaStream_t)stream : db->stream;
    const uint64_t wpq = (((n_rows + 31) >> 5) + 1 + 3) & ~3ull;  // (<= kNoAllocWords by the check above)
    CU(cudaMemsetAsync(ctx.noalloc_bits, 0, wpq * 4, st));         // the array keeps the last call's bits: start from zero
    AnswerPlan ap{};
    ap.db = db->d_slots;
    ap.row_bytes = 16;  // unused in BITS mode
    fill_answer_split(db, geom, &ap);
    ap.bits_out = ctx.noalloc_bits;
    CU(launch_answer(qc, ap, ctx.d_t0, ctx.smem_base, ctx.num_sms, st));
    CU(launch_unpack_bits(ctx.noalloc_bits, d_bits_out, n_rows, st));
    return PIR_OK;
}

extern "C" int pir_answer_with_bits(pir_db* db, uint32_t group_size, const uint8_t* bits, uint64_t n_rows,
                                    void* out, size_t out_len) {
    if (!db || (!bits && n_rows) || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    uint64_t first_row, rows;
    int rc = shard_rows(db, group_size, &first_row, &rows);
    if (rc) return rc;
    if (n_rows != rows) return fail(PIR_ERR_BAD_ARG, "n_rows %llu != shard rows %llu", (unsigned long long)n_rows,
                                    (unsigned long long)rows);
    const size_t ans = (size_t)group_size * db->slot_bytes;
    if (out_len < ans) return fail(PIR_ERR_BAD_ARG, "out_len %zu < %zu", out_len, ans);
    const uint64_t row_bytes = (uint64_t)group_size * db->pitch;
    if ((rc = ensure_out(db, row_bytes))) return rc;
    if ((rc = ensure_bits(db, n_rows ? n_rows : 1))) return rc;
    CU(cudaMemsetAsync(db->d_out, 0, row_bytes, db->stream));
    if (n_rows) {
        CU(cudaMemcpyAsync(db->d_bits, bits, n_rows, cudaMemcpyHostToDevice, db->stream));
        CU(launch_scan_bytes(db->d_slots, row_bytes, n_rows, db->d_bits, (uint32_t*)db->d_out, db->ctx->num_sms,
                             db->stream));
    }
    CU(cudaMemcpyAsync(db->h_out, db->d_out, row_bytes, cudaMemcpyDeviceToHost, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    compact_answer(db, group_size, db->h_out, (uint8_t*)out);
    return PIR_OK;
}

// Scratch of pir_dpf_eval_points, one per device, grow-only: the binding of (*Dpf).Evaluate2P (one point per
// call in the reference, dpf/server.go:55) must not pay a cudaMalloc/cudaFree per call.
namespace {
struct EvalScratch {
    std::mutex mu;
    cudaStream_t stream = nullptr;
    uint64_t* d_x = nullptr;
    int64_t* d_v = nullptr;
    uint8_t* h_pin = nullptr;  // pinned staging: n points in, n values out
    uint64_t cap = 0;
};
EvalScratch g_eval[64];

int eval_scratch_reserve(EvalScratch& sc, uint64_t n) {
    if (!sc.stream) CU(cudaStreamCreateWithFlags(&sc.stream, cudaStreamNonBlocking));
    if (n <= sc.cap) return PIR_OK;
    uint64_t cap = sc.cap ? sc.cap : 1024;
    while (cap < n) cap *= 2;
    if (sc.d_x) cudaFree(sc.d_x);
    if (sc.d_v) cudaFree(sc.d_v);
    if (sc.h_pin) cudaFreeHost(sc.h_pin);
    sc.d_x = nullptr, sc.d_v = nullptr, sc.h_pin = nullptr, sc.cap = 0;
    CU(cudaMalloc(&sc.d_x, cap * 8));
    CU(cudaMalloc(&sc.d_v, cap * 8));
    CU(cudaMallocHost(&sc.h_pin, cap * 16));
    sc.cap = cap;
    return PIR_OK;
}
}  // namespace

extern "C" int pir_dpf_eval_points(int device, const pir_query* q, const uint64_t* xs, uint64_t n, int64_t* out) {
    if (!q || (!xs && n) || (!out && n)) return fail(PIR_ERR_BAD_ARG, "null argument");
    DeviceCtx* ctx = nullptr;
    int rc = device_ctx(device, &ctx);
    if (rc) return rc;
    QueryConst qc;
    if ((rc = build_query_const(q, q->num_bits, &qc))) return rc;
    if (n == 0) return PIR_OK;
    DeviceGuard g(device);
    EvalScratch& sc = g_eval[device];
    std::lock_guard<std::mutex> lk(sc.mu);
    if ((rc = eval_scratch_reserve(sc, n))) return rc;
    memcpy(sc.h_pin, xs, n * 8);
    CU(cudaMemcpyAsync(sc.d_x, sc.h_pin, n * 8, cudaMemcpyHostToDevice, sc.stream));
    CU(launch_keyword_walk(qc, sc.d_x, 0, n, sc.d_v, nullptr, ctx->d_t0, ctx->smem_base, ctx->num_sms, sc.stream));
    CU(cudaMemcpyAsync(sc.h_pin + sc.cap * 8, sc.d_v, n * 8, cudaMemcpyDeviceToHost, sc.stream));
    CU(cudaStreamSynchronize(sc.stream));
    memcpy(out, sc.h_pin + sc.cap * 8, n * 8);
    return PIR_OK;
}

// ---- client side (row N1 of the scope table): key generation on the host ------------------------
namespace {
// AES-128 encryption of one block on the host, straight from FIPS-197 (byte oriented).
void aes_encrypt_host(const uint32_t rk[44], const uint8_t in[16], uint8_t out[16]) {
    const AesHost& A = aes_host();
    uint8_t st[16], tmp[16];
    auto add_key = [&](int round) {
        for (int c = 0; c < 4; c++) {
            uint32_t w = rk[4 * round + c];
            for (int r = 0; r < 4; r++) st[4 * c + r] ^= (uint8_t)(w >> (8 * r));
        }
    };
    auto xt = [](uint8_t v) { return (uint8_t)((v << 1) ^ ((v & 0x80) ? 0x1b : 0)); };
    memcpy(st, in, 16);
    add_key(0);
    for (int round = 1; round <= 10; round++) {
        for (int c = 0; c < 4; c++)
            for (int r = 0; r < 4; r++) tmp[4 * c + r] = A.sbox[st[4 * ((c + r) & 3) + r]];
        if (round < 10) {
            for (int c = 0; c < 4; c++) {
                const uint8_t* a = tmp + 4 * c;
                uint8_t all = (uint8_t)(a[0] ^ a[1] ^ a[2] ^ a[3]);
                for (int r = 0; r < 4; r++) st[4 * c + r] = (uint8_t)(a[r] ^ all ^ xt((uint8_t)(a[r] ^ a[(r + 1) & 3])));
            }
        } else {
            memcpy(st, tmp, 16);
        }
        add_key(round);
    }
    memcpy(out, st, 16);
}

// prf(x, FixedBlocks, 3, ...) of dpf/common.go:62-75
void prf3_host(const uint32_t rk[3][44], const uint8_t x[16], uint8_t out[48]) {
    for (int k = 0; k < 3; k++) {
        aes_encrypt_host(rk[k], x, out + 16 * k);
        for (int j = 0; j < 16; j++) out[16 * k + j] ^= x[j];
    }
}

// binary.Varint(buf[:8]) as used at dpf/client.go:141-142
int64_t varint8_host(const uint8_t* s) {
    uint64_t acc = 0;
    for (int j = 0; j < 8; j++) {
        if (s[j] < 0x80) {
            uint64_t ux = acc | ((uint64_t)s[j] << (7 * j));
            uint64_t v = ux >> 1;
            return (int64_t)((ux & 1) ? ~v : v);
        }
        acc |= (uint64_t)(s[j] & 0x7f) << (7 * j);
    }
    return 0;
}
}  // namespace

// dpf.ClientInitialize + (*Dpf).GenerateTwoServer (dpf/client.go:17-150).  `tape` supplies the 97
// bytes the reference draws from crypto/rand, in its order (4x16 PRF keys, 17, 16); NULL = /dev/urandom.
extern "C" int pir_client_gen_two_server(uint32_t num_bits, uint64_t a, uint64_t b, const uint8_t* tape,
                                         pir_query* q0, pir_query* q1) {
    if (!q0 || !q1) return fail(PIR_ERR_BAD_ARG, "null key output");
    if (num_bits == 0 || num_bits > PIR_MAX_NUM_BITS) return fail(PIR_ERR_BAD_KEY, "num_bits %u out of range", num_bits);
    uint8_t rnd[97];
    if (tape) {
        memcpy(rnd, tape, 97);
    } else {
        FILE* f = fopen("/dev/urandom", "rb");
        if (!f || fread(rnd, 1, 97, f) != 97) {
            if (f) fclose(f);
            return fail(PIR_ERR_BAD_ARG, "cannot read /dev/urandom");
        }
        fclose(f);
    }
    pir_query* qs[2] = {q0, q1};
    for (int p = 0; p < 2; p++) {
        memset(qs[p], 0, sizeof(pir_query));
        memcpy(qs[p]->prf_keys, rnd, 64);
        qs[p]->share_number = (uint8_t)p;
        qs[p]->is_two_party = 1;
        qs[p]->num_bits = num_bits;
        qs[p]->group_size = 1;
    }
    uint32_t rk[3][44];
    for (int k = 0; k < 3; k++) aes_host().expand(rnd + 16 * k, rk[k]);
    memcpy(q0->s_init, rnd + 64, 16);
    q0->t_init = rnd[80] & 1;
    memcpy(q1->s_init, rnd + 81, 16);
    q1->t_init = q0->t_init ^ 1;
    uint8_t s[2][16], t[2] = {q0->t_init, q1->t_init};
    memcpy(s[0], q0->s_init, 16);
    memcpy(s[1], q1->s_init, 16);
    for (uint32_t i = 0; i < num_bits; i++) {
        uint8_t g[2][48];
        prf3_host(rk, s[0], g[0]);
        prf3_host(rk, s[1], g[1]);
        const uint8_t abit = (uint8_t)((a >> (num_bits - 1 - i)) & 1);
        const int keep = abit ? 17 : 0, lose = abit ? 0 : 17;
        uint8_t* cw = q0->cw[i];
        for (int j = 0; j < 16; j++) cw[j] = g[0][lose + j] ^ g[1][lose + j];
        cw[16] = (uint8_t)((g[0][16] & 1) ^ (g[1][16] & 1) ^ abit ^ 1);
        cw[17] = (uint8_t)((g[0][33] & 1) ^ (g[1][33] & 1) ^ abit);
        const uint8_t tkeep = abit ? cw[17] : cw[16];
        for (int p = 0; p < 2; p++) {
            for (int j = 0; j < 16; j++) s[p][j] = (uint8_t)(g[p][keep + j] ^ (t[p] ? cw[j] : 0));
            t[p] = (uint8_t)((g[p][keep + 16] & 1) ^ (tkeep & t[p]));
        }
        memcpy(q1->cw[i], cw, PIR_CW_BYTES);
    }
    uint64_t f = b - (uint64_t)varint8_host(s[0]) + (uint64_t)varint8_host(s[1]);
    if (t[1] == 1) f = 0 - f;
    q0->final_cw = q1->final_cw = (int64_t)f;
    return PIR_OK;
}

// ---- debugging hooks for the CPU test-suite (host logic only, no device needed) -----------------
extern "C" void pir_debug_round_keys(const uint8_t key[16], uint32_t rk[44]) { aes_host().expand(key, rk); }
extern "C" void pir_debug_t0(uint32_t t0[256]) { memcpy(t0, aes_host().t0, 1024); }
extern "C" void pir_debug_plan(uint32_t nb, uint64_t first_row, uint64_t n_rows, int num_sms, uint32_t* D,
                               uint32_t* B, uint32_t* Fc, uint64_t* task_first, uint64_t* task_count) {
    Plan p = make_plan(nb, first_row, n_rows, num_sms);
    *D = p.D; *B = p.B; *Fc = p.Fc; *task_first = p.task_first; *task_count = p.task_count;
}

// ---- secret-shared ASPIR audit (aspir.go:227-295) -------------------------------------------------
extern "C" int pir_audit_shared_query(pir_db* key_db, const pir_query* q, const void* auth_token, size_t token_len,
                                      void* out, size_t out_len) {
    if (!key_db || !q || (!auth_token && token_len) || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (q->group_size != 1)  // aspir.go:275-277: len(res.Shares) = query.GroupSize must be 1
        return fail(PIR_ERR_BAD_ARG, "Invalid challenge ciphertext result (query group size %u, the key database has 1)",
                    q->group_size);
    if (out_len < key_db->slot_bytes) return fail(PIR_ERR_BAD_ARG, "out_len %zu < %u", out_len, key_db->slot_bytes);
    int rc = pir_answer(key_db, q, out, out_len);
    if (rc) return rc;
    const size_t n = token_len < key_db->slot_bytes ? token_len : key_db->slot_bytes;
    pir_xor_combine(out, auth_token, n);
    return PIR_OK;
}

extern "C" int pir_audit_shared_query_with_bits(pir_db* key_db, uint32_t group_size, const uint8_t* bits, uint64_t n_rows,
                                                const void* auth_token, size_t token_len, void* out, size_t out_len) {
    if (!key_db || (!auth_token && token_len) || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (group_size != 1)
        return fail(PIR_ERR_BAD_ARG, "Invalid challenge ciphertext result (query group size %u, the key database has 1)",
                    group_size);
    int rc = pir_answer_with_bits(key_db, 1, bits, n_rows, out, out_len);
    if (rc) return rc;
    const size_t n = token_len < key_db->slot_bytes ? token_len : key_db->slot_bytes;
    pir_xor_combine(out, auth_token, n);
    return PIR_OK;
}

// ---- multi-GPU combine ------------------------------------------------------------
// Layout of a combine buffer on the owner GPU: [n_slots x slot_stride answer bytes][n_slots x u32 arrivals]
// [u32 timeout flag].  slot_stride is a multiple of 16.
struct pir_combine {
    int device = 0;          // device of THIS process that uses the mapping
    bool owner = false;
    uint32_t n_slots = 0;
    size_t stride = 0;
    uint8_t* base = nullptr;  // device pointer valid in this process
    cudaStream_t stream = nullptr;  // owner: reset / wait / read
    uint8_t* h_buf = nullptr;       // owner: pinned staging of one slot + the timeout flag
    uint8_t* h_buf_dev = nullptr;   // the same buffer as the device sees it (mapped), or null
    int clock_khz = 0;
    std::mutex mu;
    uint32_t* arrivals(uint32_t slot) const { return (uint32_t*)(base + (size_t)n_slots * stride) + slot; }
    uint32_t* timeout_flag() const { return (uint32_t*)(base + (size_t)n_slots * stride) + n_slots; }
    size_t total_bytes() const { return (size_t)n_slots * stride + ((size_t)n_slots + 1) * 4; }
};

static int combine_args(uint32_t n_slots, size_t slot_stride) {
    if (n_slots == 0 || slot_stride == 0 || (slot_stride & 15))
        return fail(PIR_ERR_BAD_ARG, "combine needs n_slots >= 1 and a slot_stride that is a non-zero multiple of 16");
    return PIR_OK;
}

extern "C" int pir_combine_create(int device, uint32_t n_slots, size_t slot_stride, pir_combine** out) {
    if (!out) return fail(PIR_ERR_BAD_ARG, "out is null");
    *out = nullptr;
    int rc = combine_args(n_slots, slot_stride);
    if (rc) return rc;
    DeviceCtx* ctx = nullptr;
    if ((rc = device_ctx(device, &ctx))) return rc;
    DeviceGuard g(device);
    pir_combine* c = new (std::nothrow) pir_combine();
    if (!c) return fail(PIR_ERR_OOM, "host allocation failed");
    c->device = device;
    c->owner = true;
    c->n_slots = n_slots;
    c->stride = slot_stride;
    cudaError_t e = cudaMalloc(&c->base, c->total_bytes());  // plain cudaMalloc: exportable with CUDA IPC
    if (e == cudaSuccess) e = cudaMemset(c->base, 0, c->total_bytes());
    if (e == cudaSuccess) e = cudaDeviceSynchronize();  // (the slots are used from non-blocking streams and other processes)
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMallocHost(&c->h_buf, slot_stride + 16);
    if (e == cudaSuccess) {
        memset(c->h_buf, 0, slot_stride + 16);
        void* dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, c->h_buf, 0) == cudaSuccess) c->h_buf_dev = (uint8_t*)dp;
        else cudaGetLastError();
    }
    c->clock_khz = ctx->clock_khz;
    if (e != cudaSuccess) {
        int code = e == cudaErrorMemoryAllocation ? PIR_ERR_OOM : PIR_ERR_CUDA;
        fail(code, "pir_combine_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        pir_combine_destroy(c);
        return code;
    }
    *out = c;
    return PIR_OK;
}

extern "C" int pir_combine_export(pir_combine* c, uint8_t handle[PIR_IPC_HANDLE_BYTES]) {
    if (!c || !handle) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (!c->owner) return fail(PIR_ERR_BAD_ARG, "only the owner can export a combine buffer");
    static_assert(sizeof(cudaIpcMemHandle_t) == PIR_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    DeviceGuard g(c->device);
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, c->base));
    memcpy(handle, &h, sizeof h);
    return PIR_OK;
}

extern "C" int pir_combine_open(int device, const uint8_t handle[PIR_IPC_HANDLE_BYTES], uint32_t n_slots,
                                size_t slot_stride, pir_combine** out) {
    if (!out || !handle) return fail(PIR_ERR_BAD_ARG, "null argument");
    *out = nullptr;
    int rc = combine_args(n_slots, slot_stride);
    if (rc) return rc;
    DeviceCtx* ctx = nullptr;
    if ((rc = device_ctx(device, &ctx))) return rc;
    DeviceGuard g(device);
    pir_combine* c = new (std::nothrow) pir_combine();
    if (!c) return fail(PIR_ERR_OOM, "host allocation failed");
    c->device = device;
    c->owner = false;
    c->n_slots = n_slots;
    c->stride = slot_stride;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        delete c;
        cudaGetLastError();
        return fail(PIR_ERR_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
    }
    c->base = (uint8_t*)p;
    *out = c;
    return PIR_OK;
}

extern "C" int pir_combine_destroy(pir_combine* c) {
    if (!c) return PIR_OK;
    DeviceGuard g(c->device);
    if (c->owner) {
        if (c->stream) {
            cudaStreamSynchronize(c->stream);
            cudaStreamDestroy(c->stream);
        }
        if (c->h_buf) cudaFreeHost(c->h_buf);
        if (c->base) cudaFree(c->base);
    } else if (c->base) {
        cudaDeviceSynchronize();
        cudaIpcCloseMemHandle(c->base);
    }
    delete c;
    return PIR_OK;
}

extern "C" int pir_combine_reset(pir_combine* c, uint32_t first, uint32_t count) {
    if (!c) return fail(PIR_ERR_BAD_ARG, "null handle");
    if (!c->owner) return fail(PIR_ERR_BAD_ARG, "only the owner resets combine slots");
    if (first > c->n_slots || count > c->n_slots - first) return fail(PIR_ERR_BAD_ARG, "slot range out of bounds");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    CU(cudaMemsetAsync(c->base + (size_t)first * c->stride, 0, (size_t)count * c->stride, c->stream));
    CU(cudaMemsetAsync(c->arrivals(first), 0, (size_t)count * 4, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return PIR_OK;
}

extern "C" void* pir_combine_slot_ptr(pir_combine* c, uint32_t slot) {
    if (!c || slot >= c->n_slots) return nullptr;
    return c->base + (size_t)slot * c->stride;
}

extern "C" int pir_answer_accumulate_device(pir_db* db, const pir_query* q, pir_combine* c, uint32_t slot,
                                            void* stream) {
    if (!db || !q || !c) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (slot >= c->n_slots) return fail(PIR_ERR_BAD_ARG, "slot %u out of range (%u slots)", slot, c->n_slots);
    if (c->device != db->device)
        return fail(PIR_ERR_BAD_ARG, "combine buffer is mapped on device %d, the shard lives on device %d", c->device,
                    db->device);
    std::lock_guard<std::mutex> lk(db->mu);
    DeviceGuard g(db->device);
    cudaStream_t st = stream ? (cudaStream_t)stream : db->stream;
    QueryGeom geom;
    int rc = query_geom(db, q, &geom);
    if (rc) return rc;
    if (geom.row_bytes > c->stride) return fail(PIR_ERR_BAD_ARG, "answer of %llu bytes exceeds the slot stride %zu",
                                                (unsigned long long)geom.row_bytes, c->stride);
    bool signalled = false;
    uint8_t* d_slot = c->base + (size_t)slot * c->stride;
    // The CTAs fold into the handle's local accumulator and the launch's LAST CTA forwards the shard's sum to the slot
    // (possibly on a peer GPU) and signals: a CTA's exit does not wait for a round trip over NVLink, so the next launch's
    // CTA gets the SM earlier.  (direct_combine: every CTA XORs into the slot itself, as in round 1.)
    if (!db->tune.direct_combine && !db->tune.signal_kernel && single_launch_ok(db, q, geom))
        rc = enqueue_answer(db, q, geom, db->d_acc, st, /*zero_first=*/false, c->arrivals(slot), &signalled, false, (uint32_t*)d_slot);
    else
        rc = enqueue_answer(db, q, geom, d_slot, st, /*zero_first=*/false, c->arrivals(slot), &signalled);
    if (rc) return rc;
    if (!signalled) CU(launch_signal_arrival(c->arrivals(slot), st));
    return PIR_OK;
}

extern "C" int pir_combine_wait_read(pir_combine* c, uint32_t slot, uint32_t expected, void* out, size_t out_len,
                                     uint32_t timeout_ms) {
    if (!c || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (!c->owner) return fail(PIR_ERR_BAD_ARG, "only the owner reads combine slots");
    if (slot >= c->n_slots || out_len > c->stride) return fail(PIR_ERR_BAD_ARG, "slot or length out of range");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    const uint64_t cycles = (uint64_t)(c->clock_khz > 0 ? c->clock_khz : 1000000) * (timeout_ms ? timeout_ms : 1);
    CU(launch_wait_arrivals(c->arrivals(slot), expected, cycles, c->timeout_flag(), c->stream));
    CU(cudaMemcpyAsync(c->h_buf, c->base + (size_t)slot * c->stride, out_len, cudaMemcpyDeviceToHost, c->stream));
    uint32_t* h_flag = (uint32_t*)(c->h_buf + c->stride);
    CU(cudaMemcpyAsync(h_flag, c->timeout_flag(), 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (*h_flag) {
        cudaMemsetAsync(c->timeout_flag(), 0, 4, c->stream);
        cudaStreamSynchronize(c->stream);
        return fail(PIR_ERR_CUDA, "timed out after %u ms waiting for %u shard arrivals at combine slot %u", timeout_ms,
                    expected, slot);
    }
    memcpy(out, c->h_buf, out_len);
    return PIR_OK;
}

// Owner only: enqueue the wait for `expected` arrivals at `slot` on `stream` (device-side, bounded) and return.
extern "C" int pir_combine_wait_device(pir_combine* c, uint32_t slot, uint32_t expected, uint32_t timeout_ms, void* stream) {
    if (!c) return fail(PIR_ERR_BAD_ARG, "null handle");
    if (!c->owner) return fail(PIR_ERR_BAD_ARG, "only the owner waits on combine slots");
    if (slot >= c->n_slots) return fail(PIR_ERR_BAD_ARG, "slot out of range");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    const uint64_t cycles = (uint64_t)(c->clock_khz > 0 ? c->clock_khz : 1000000) * (timeout_ms ? timeout_ms : 1);
    CU(launch_wait_arrivals(c->arrivals(slot), expected, cycles, c->timeout_flag(), stream ? (cudaStream_t)stream : c->stream));
    return PIR_OK;
}

// Owner only: pir_combine_wait_read, then the slot and its arrival counter are zeroed behind the copy on the same
// stream, with ONE synchronisation: on return the slot is free for the next query.
extern "C" int pir_combine_wait_read_reset(pir_combine* c, uint32_t slot, uint32_t expected, void* out, size_t out_len,
                                           uint32_t timeout_ms) {
    if (!c || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (!c->owner) return fail(PIR_ERR_BAD_ARG, "only the owner reads combine slots");
    if (slot >= c->n_slots || out_len > c->stride) return fail(PIR_ERR_BAD_ARG, "slot or length out of range");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    const uint64_t cycles = (uint64_t)(c->clock_khz > 0 ? c->clock_khz : 1000000) * (timeout_ms ? timeout_ms : 1);
    uint32_t* h_flag = (uint32_t*)(c->h_buf + c->stride);
    if (c->h_buf_dev) {
        // one launch: wait for the arrivals, move the answer to the mapped pinned buffer, zero slot and counter
        CU(launch_wait_read_reset(c->arrivals(slot), expected, cycles, (uint32_t*)(c->base + (size_t)slot * c->stride),
                                  (uint32_t*)c->h_buf_dev, (uint32_t*)(c->h_buf_dev + c->stride), (uint32_t)(c->stride / 4), c->stream));
        CU(cudaStreamSynchronize(c->stream));
        if (*h_flag) return fail(PIR_ERR_CUDA, "timed out after %u ms waiting for %u shard arrivals at combine slot %u", timeout_ms,
                                 expected, slot);
        memcpy(out, c->h_buf, out_len);
        return PIR_OK;
    }
    CU(launch_wait_arrivals(c->arrivals(slot), expected, cycles, c->timeout_flag(), c->stream));
    CU(cudaMemcpyAsync(c->h_buf, c->base + (size_t)slot * c->stride, out_len, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(h_flag, c->timeout_flag(), 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemsetAsync(c->base + (size_t)slot * c->stride, 0, c->stride, c->stream));
    CU(cudaMemsetAsync(c->arrivals(slot), 0, 4, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (*h_flag) {
        cudaMemsetAsync(c->timeout_flag(), 0, 4, c->stream);
        cudaStreamSynchronize(c->stream);
        return fail(PIR_ERR_CUDA, "timed out after %u ms waiting for %u shard arrivals at combine slot %u", timeout_ms,
                    expected, slot);
    }
    memcpy(out, c->h_buf, out_len);
    return PIR_OK;
}

// ---- pir_server: one process, all shards -----------------------------------------------------------
struct pir_server {
    std::vector<pir_db*> shards;
    std::vector<cudaEvent_t> done;   // per shard: its partial has been accumulated
    cudaEvent_t zeroed = nullptr;    // owner stream: the accumulator has been cleared
    int owner_device = 0;
    uint64_t db_size = 0;
    uint32_t slot_bytes = 0, pitch = 0;
    uint32_t shard_align = 1;
    uint8_t* d_acc = nullptr;        // on the owner device; peers reach it over NVLink (peer access)
    uint64_t acc_cap = 0;
    uint8_t* h_acc = nullptr;        // pinned
    cudaStream_t stream = nullptr;   // owner stream
    std::mutex mu;
    // Single index queries (pir_server_answer): a ring of kHostRing combine slots on the owner device, each with an arrival
    // counter, a pinned host slot and an event.  Every shard's kernel XORs into slot k and its last CTA bumps the counter
    // (one launch per shard, no memset, no per-shard event); ONE kernel on the owner waits for the arrivals, moves the
    // answer to the pinned slot and leaves slot and counter zeroed.  A caller enqueues under the server's mutex and waits
    // for its own event with the mutex released, so queries of several threads are in flight together.
    uint8_t* d_ring = nullptr;       // [kHostRing][kAnswerMaxRowBytes] + [kHostRing] u32 arrivals, zero between queries
    uint8_t* h_ring = nullptr;       // pinned, mapped: [kHostRing][kAnswerMaxRowBytes + 16] (answer + time-out flag)
    uint8_t* h_ring_dev = nullptr;
    cudaEvent_t ring_ev[kHostRing] = {};
    uint32_t ring_busy = 0;
    std::condition_variable ring_cv;
    int clock_khz = 0;
    uint32_t* ring_arrivals(uint32_t k) const { return (uint32_t*)(d_ring + (size_t)kHostRing * kAnswerMaxRowBytes) + k; }
};

namespace {
int server_ensure_acc(pir_server* s, uint64_t n) {
    if (n <= s->acc_cap) return PIR_OK;
    DeviceGuard g(s->owner_device);
    if (s->d_acc) cudaFree(s->d_acc);
    if (s->h_acc) cudaFreeHost(s->h_acc);
    s->d_acc = nullptr;
    s->h_acc = nullptr;
    s->acc_cap = 0;
    CU(cudaMalloc(&s->d_acc, n));
    CU(cudaMallocHost(&s->h_acc, n));
    s->acc_cap = n;
    return PIR_OK;
}
}  // namespace

extern "C" int pir_server_create(const int* devices, int n_devices, uint64_t db_size, uint32_t slot_bytes,
                                 uint32_t shard_align, pir_server** out) {
    if (!out) return fail(PIR_ERR_BAD_ARG, "out is null");
    *out = nullptr;
    if (!devices || n_devices <= 0) return fail(PIR_ERR_BAD_ARG, "need at least one device");
    if (db_size == 0 || slot_bytes == 0) return fail(PIR_ERR_BAD_ARG, "db_size and slot_bytes must be non-zero");
    if (shard_align == 0) shard_align = 1;
    pir_server* s = new (std::nothrow) pir_server();
    if (!s) return fail(PIR_ERR_OOM, "host allocation failed");
    s->db_size = db_size;
    s->slot_bytes = slot_bytes;
    s->pitch = (slot_bytes + 15u) & ~15u;
    s->owner_device = devices[0];
    s->shard_align = shard_align;
    const uint64_t units = db_size / shard_align;  // whole alignment units; the ragged tail goes to the last shard
    int rc = PIR_OK;
    for (int i = 0; i < n_devices && rc == PIR_OK; i++) {
        const uint64_t b = units * (uint64_t)i / (uint64_t)n_devices * shard_align;
        const uint64_t e = i == n_devices - 1 ? db_size : units * (uint64_t)(i + 1) / (uint64_t)n_devices * shard_align;
        pir_db* db = nullptr;
        rc = pir_db_create(devices[i], db_size, slot_bytes, b, e - b, &db);
        if (rc == PIR_OK) s->shards.push_back(db);
    }
    // peer access: every shard's device must be able to store into the owner's memory
    for (size_t i = 0; i < s->shards.size() && rc == PIR_OK; i++) {
        const int dev = s->shards[i]->device;
        if (dev == s->owner_device) continue;
        int can = 0;
        cudaError_t e = cudaDeviceCanAccessPeer(&can, dev, s->owner_device);
        if (e != cudaSuccess || !can) {
            rc = fail(PIR_ERR_CUDA, "device %d cannot access device %d as a peer (NVLink/P2P needed for the combine)", dev,
                      s->owner_device);
            break;
        }
        DeviceGuard g(dev);
        e = cudaDeviceEnablePeerAccess(s->owner_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) {
            cudaGetLastError();
            e = cudaSuccess;
        }
        if (e != cudaSuccess) rc = fail(PIR_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", dev, s->owner_device,
                                        cudaGetErrorString(e));
    }
    if (rc == PIR_OK) {
        DeviceGuard g(s->owner_device);
        cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->zeroed, cudaEventDisableTiming);
        const size_t ring_bytes = (size_t)kHostRing * kAnswerMaxRowBytes + kHostRing * 4;
        if (e == cudaSuccess) e = cudaMalloc(&s->d_ring, ring_bytes);
        if (e == cudaSuccess) e = cudaMemsetAsync(s->d_ring, 0, ring_bytes, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e == cudaSuccess) e = cudaHostAlloc(&s->h_ring, kHostRing * (kAnswerMaxRowBytes + 16), cudaHostAllocMapped);
        if (e == cudaSuccess) {
            memset(s->h_ring, 0, kHostRing * (kAnswerMaxRowBytes + 16));
            void* dp = nullptr;
            if (cudaHostGetDevicePointer(&dp, s->h_ring, 0) == cudaSuccess) s->h_ring_dev = (uint8_t*)dp;
            else cudaGetLastError();
        }
        for (uint32_t k = 0; k < kHostRing && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&s->ring_ev[k], cudaEventDisableTiming);
        if (e == cudaSuccess && !s->shards.empty()) s->clock_khz = s->shards[0]->ctx->clock_khz;
        if (e != cudaSuccess) rc = fail(PIR_ERR_CUDA, "pir_server_create: %s", cudaGetErrorString(e));
    }
    for (size_t i = 0; i < s->shards.size() && rc == PIR_OK; i++) {
        DeviceGuard g(s->shards[i]->device);
        cudaEvent_t ev = nullptr;
        cudaError_t e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        if (e != cudaSuccess) rc = fail(PIR_ERR_CUDA, "pir_server_create: %s", cudaGetErrorString(e));
        else s->done.push_back(ev);
    }
    if (rc != PIR_OK) {
        char keep[sizeof tl_err];
        memcpy(keep, tl_err, sizeof keep);
        pir_server_destroy(s);
        memcpy(tl_err, keep, sizeof keep);
        return rc;
    }
    *out = s;
    return PIR_OK;
}

extern "C" int pir_server_destroy(pir_server* s) {
    if (!s) return PIR_OK;
    for (size_t i = 0; i < s->done.size(); i++) {
        DeviceGuard g(s->shards[i]->device);
        cudaEventDestroy(s->done[i]);
    }
    for (pir_db* db : s->shards) pir_db_destroy(db);
    {
        DeviceGuard g(s->owner_device);
        if (s->stream) {
            cudaStreamSynchronize(s->stream);
            cudaStreamDestroy(s->stream);
        }
        if (s->zeroed) cudaEventDestroy(s->zeroed);
        if (s->d_ring) cudaFree(s->d_ring);
        if (s->h_ring) cudaFreeHost(s->h_ring);
        for (uint32_t k = 0; k < kHostRing; k++)
            if (s->ring_ev[k]) cudaEventDestroy(s->ring_ev[k]);
        if (s->d_acc) cudaFree(s->d_acc);
        if (s->h_acc) cudaFreeHost(s->h_acc);
    }
    delete s;
    return PIR_OK;
}

extern "C" int pir_server_num_shards(const pir_server* s) { return s ? (int)s->shards.size() : 0; }

extern "C" int pir_server_shard(pir_server* s, int i, pir_db** db) {
    if (!s || !db) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (i < 0 || (size_t)i >= s->shards.size()) return fail(PIR_ERR_BAD_ARG, "shard %d out of range", i);
    *db = s->shards[(size_t)i];
    return PIR_OK;
}

extern "C" int pir_server_set_option(pir_server* s, const char* name, int64_t value) {
    if (!s || !name) return fail(PIR_ERR_BAD_ARG, "null argument");
    for (pir_db* db : s->shards) {
        int rc = pir_db_set_option(db, name, value);
        if (rc) return rc;
    }
    return PIR_OK;
}

extern "C" int pir_server_upload(pir_server* s, uint64_t slot_index, const void* host_slots, uint64_t n_slots) {
    if (!s || (!host_slots && n_slots)) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (slot_index > s->db_size || n_slots > s->db_size - slot_index)
        return fail(PIR_ERR_BAD_ARG, "slots [%llu, +%llu) exceed db_size", (unsigned long long)slot_index,
                    (unsigned long long)n_slots);
    for (pir_db* db : s->shards) {
        const uint64_t b = slot_index > db->slot_begin ? slot_index : db->slot_begin;
        const uint64_t e0 = slot_index + n_slots, e1 = db->slot_begin + db->slot_count;
        const uint64_t e = e0 < e1 ? e0 : e1;
        if (b >= e) continue;
        int rc = pir_db_upload(db, b, (const uint8_t*)host_slots + (b - slot_index) * (uint64_t)s->slot_bytes, e - b);
        if (rc) return rc;
    }
    return PIR_OK;
}

extern "C" int pir_server_set_keywords(pir_server* s, const uint64_t* keywords, uint64_t n) {
    if (!s || !keywords) return fail(PIR_ERR_BAD_ARG, "null argument");
    if (n != s->db_size)
        return fail(PIR_ERR_BAD_ARG, "expected %llu keywords (one per slot), got %llu", (unsigned long long)s->db_size,
                    (unsigned long long)n);
    for (pir_db* db : s->shards) {  // every shard keeps the keywords of its own slots only
        int rc = pir_db_set_keywords(db, keywords + db->slot_begin, db->slot_count);
        if (rc) return rc;
        // ... and, for keyword queries at the group size the shards are aligned to, those of its own ROWS: row r evaluates
        // Keywords[r] (db.go:132-135), which for W > 1 belongs to an earlier shard's slots
        if (s->shard_align > 1) {
            uint64_t first = 0, rows = 0;
            if ((rc = pir_db_rows(db, s->shard_align, &first, &rows))) return rc;
            if ((rc = pir_db_set_row_keywords(db, s->shard_align, keywords + first, rows))) return rc;
        }
    }
    return PIR_OK;
}

extern "C" int pir_server_fill_synthetic(pir_server* s, uint64_t seed) {
    if (!s) return fail(PIR_ERR_BAD_ARG, "null handle");
    for (pir_db* db : s->shards) {
        int rc = pir_db_fill_synthetic(db, seed);
        if (rc) return rc;
    }
    return PIR_OK;
}

extern "C" int pir_server_answer_batch(pir_server* s, const pir_query* qs, uint32_t n_queries, void* out,
                                       size_t out_len) {
    if (!s || !qs || !out || n_queries == 0) return fail(PIR_ERR_BAD_ARG, "null or empty argument");
    std::lock_guard<std::mutex> lk(s->mu);
    const uint32_t W = qs[0].group_size;
    if (W == 0) return fail(PIR_ERR_BAD_ARG, "group_size must be >= 1");
    const uint64_t row_bytes = (uint64_t)W * s->pitch;
    const size_t ans = (size_t)W * s->slot_bytes;
    if (out_len < ans * n_queries) return fail(PIR_ERR_BAD_ARG, "out_len %zu < %zu", out_len, ans * n_queries);
    int rc = server_ensure_acc(s, row_bytes * n_queries);
    if (rc) return rc;
    // Validate every query against every shard BEFORE the first memset or launch: an error found at shard i > 0
    // (a group size that does not divide that shard's first slot, a bad key) must not leave the kernels of the shards
    // before it XOR-ing into the accumulator while the caller already sees a failure.
    std::vector<QueryGeom> geoms(s->shards.size());
    std::vector<QueryConst> qcs(n_queries);  // marshalled once, shared by all shards (the key material is the same)
    for (size_t i = 0; i < s->shards.size(); i++) {
        for (uint32_t k = 0; k < n_queries; k++) {
            if (qs[k].group_size != W || qs[k].is_keyword != qs[0].is_keyword)
                return fail(PIR_ERR_BAD_ARG, "all queries of a batch must share group_size and is_keyword");
            QueryGeom geom;
            if ((rc = query_geom(s->shards[i], &qs[k], &geom))) return rc;
            if (k == 0) geoms[i] = geom;
            if (i == 0 && (rc = build_query_const(&qs[k], geom.nb, &qcs[k]))) return rc;
            if (geom.nb != geoms[0].nb) return fail(PIR_ERR_BAD_ARG, "shards disagree on the domain size");
        }
        if (qs[0].is_keyword && geoms[i].n_rows) {  // (nothing has been enqueued yet)
            const uint64_t* kw = nullptr;
            if (!s->shards[i]->d_keywords)
                return fail(PIR_ERR_BAD_ARG, "keyword query on a database without keywords (pir_server_set_keywords)");
            if ((rc = shard_keywords(s->shards[i], geoms[i], &kw))) return rc;
        }
    }
    {
        DeviceGuard g(s->owner_device);
        CU(cudaMemsetAsync(s->d_acc, 0, row_bytes * n_queries, s->stream));
        CU(cudaEventRecord(s->zeroed, s->stream));
    }
    // From here on work is in flight: a failure is remembered, the remaining shards are skipped, and the streams
    // of the shards already started are joined before the call returns.
    size_t started = 0;
    for (size_t i = 0; i < s->shards.size() && rc == PIR_OK; i++) {
        pir_db* db = s->shards[i];
        std::lock_guard<std::mutex> lkdb(db->mu);
        DeviceGuard g(db->device);
        auto enqueue = [&]() -> int {
            CU(cudaStreamWaitEvent(db->stream, s->zeroed, 0));
            int r = PIR_OK;
            if (batch_uses_bits_path(db, qs, n_queries, geoms[i])) {
                // two-phase batch per shard; its scan kernel XORs into the shared accumulator like the fused one
                r = enqueue_batch_bits(db, qs, n_queries, geoms[i], s->d_acc, db->stream, /*zero_first=*/false, qcs.data());
            } else {
                for (uint32_t k = 0; k < n_queries && r == PIR_OK; k++)
                    r = enqueue_answer(db, &qs[k], geoms[i], s->d_acc + (uint64_t)k * row_bytes, db->stream, false, nullptr, nullptr,
                                       false, nullptr, &qcs[k]);
            }
            return r;
        };
        rc = enqueue();
        started = i + 1;
        if (cudaEventRecord(s->done[i], db->stream) != cudaSuccess && rc == PIR_OK)
            rc = fail(PIR_ERR_CUDA, "cudaEventRecord: %s", cudaGetErrorString(cudaGetLastError()));
    }
    {
        DeviceGuard g(s->owner_device);
        for (size_t i = 0; i < started; i++) {
            cudaError_t e = cudaStreamWaitEvent(s->stream, s->done[i], 0);
            if (e != cudaSuccess && rc == PIR_OK) rc = fail(PIR_ERR_CUDA, "cudaStreamWaitEvent: %s", cudaGetErrorString(e));
        }
        if (rc == PIR_OK) {
            cudaError_t e = cudaMemcpyAsync(s->h_acc, s->d_acc, row_bytes * n_queries, cudaMemcpyDeviceToHost, s->stream);
            if (e != cudaSuccess) rc = fail(PIR_ERR_CUDA, "cudaMemcpyAsync: %s", cudaGetErrorString(e));
        }
        cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess && rc == PIR_OK) rc = fail(PIR_ERR_CUDA, "pir_server_answer_batch: %s", cudaGetErrorString(e));
        if (rc != PIR_OK) {
            // (if an event could not be recorded the owner stream has not waited for that shard: wait for it directly)
            for (size_t i = 0; i < started; i++) {
                DeviceGuard gi(s->shards[i]->device);
                cudaStreamSynchronize(s->shards[i]->stream);
            }
            cudaGetLastError();
            return rc;
        }
    }
    for (uint32_t k = 0; k < n_queries; k++)
        for (uint32_t c = 0; c < W; c++)
            memcpy((uint8_t*)out + (size_t)k * ans + (size_t)c * s->slot_bytes,
                   s->h_acc + (uint64_t)k * row_bytes + (uint64_t)c * s->pitch, s->slot_bytes);
    return PIR_OK;
}

extern "C" int pir_server_answer(pir_server* s, const pir_query* q, void* out, size_t out_len) {
    if (!s || !q || !out) return fail(PIR_ERR_BAD_ARG, "null argument");
    // The one-launch-per-shard path (see struct pir_server) when every shard can take it; anything else -- keyword
    // queries, rows wider than 4 KB, the older kernels selected by an option -- goes through the batch path.
    std::unique_lock<std::mutex> lk(s->mu);
    const uint32_t W = q->group_size;
    bool fast = s->h_ring_dev != nullptr && W != 0 && !q->is_keyword && !dev_phase_debug_on();
    std::vector<QueryGeom> geoms(s->shards.size());
    size_t live = 0;
    for (size_t i = 0; i < s->shards.size() && fast; i++) {
        if (query_geom(s->shards[i], q, &geoms[i]) != PIR_OK) fast = false;  // (the batch path reports the error)
        else if (geoms[i].n_rows != 0) {
            live++;
            if (!single_launch_ok(s->shards[i], q, geoms[i]) || s->shards[i]->tune.signal_kernel) fast = false;
        }
    }
    if (!fast || live == 0) {
        lk.unlock();
        return pir_server_answer_batch(s, q, 1, out, out_len);
    }
    const uint64_t row_bytes = (uint64_t)W * s->pitch;
    const size_t ans = (size_t)W * s->slot_bytes;
    if (out_len < ans) return fail(PIR_ERR_BAD_ARG, "out_len %zu < %zu", out_len, ans);
    QueryConst qc;  // marshalled once, shared by all shards
    int rc = build_query_const(q, geoms[0].nb, &qc);
    if (rc) return rc;
    s->ring_cv.wait(lk, [&] { return s->ring_busy != (1u << kHostRing) - 1u; });
    uint32_t k = 0;
    while ((s->ring_busy >> k) & 1u) k++;
    s->ring_busy |= 1u << k;
    uint8_t* d_slot = s->d_ring + (size_t)k * kAnswerMaxRowBytes;
    uint8_t* h_slot = s->h_ring + (size_t)k * (kAnswerMaxRowBytes + 16);
    uint8_t* h_slot_dev = s->h_ring_dev + (size_t)k * (kAnswerMaxRowBytes + 16);
    uint32_t arrived = 0;  // shards whose arrival is (or will be) signalled
    for (size_t i = 0; i < s->shards.size() && rc == PIR_OK; i++) {
        if (geoms[i].n_rows == 0) continue;
        pir_db* db = s->shards[i];
        std::lock_guard<std::mutex> lkdb(db->mu);
        DeviceGuard g(db->device);
        bool signalled = false;
        if (!db->tune.direct_combine)  // (fold locally, the launch's last CTA forwards: see pir_answer_accumulate_device)
            rc = enqueue_answer(db, q, geoms[i], db->d_acc, db->stream, /*zero_first=*/false, s->ring_arrivals(k), &signalled, false,
                                (uint32_t*)d_slot, &qc);
        else
            rc = enqueue_answer(db, q, geoms[i], d_slot, db->stream, /*zero_first=*/false, s->ring_arrivals(k), &signalled, false, nullptr, &qc);
        if (rc == PIR_OK && signalled) arrived++;
        else if (rc == PIR_OK) rc = fail(PIR_ERR_CUDA, "shard %zu did not take the one-launch path", i);
    }
    cudaError_t e = cudaSuccess;
    {
        DeviceGuard g(s->owner_device);
        // (on a failure part-way the missing arrivals are made up for, so that the wait below ends and the slot is reset)
        for (uint32_t m = arrived; m < live && e == cudaSuccess; m++) e = launch_signal_arrival(s->ring_arrivals(k), s->stream);
        const uint64_t cycles = (uint64_t)(s->clock_khz > 0 ? s->clock_khz : 1000000) * 10000ull;  // 10 s
        if (e == cudaSuccess)
            e = launch_wait_read_reset(s->ring_arrivals(k), (uint32_t)live, cycles, (uint32_t*)d_slot, (uint32_t*)h_slot_dev,
                                       (uint32_t*)(h_slot_dev + kAnswerMaxRowBytes), (uint32_t)(row_bytes / 4), s->stream);
        if (e == cudaSuccess) e = cudaEventRecord(s->ring_ev[k], s->stream);
    }
    lk.unlock();
    if (e == cudaSuccess) e = cudaEventSynchronize(s->ring_ev[k]);
    if (e != cudaSuccess && rc == PIR_OK) rc = fail(PIR_ERR_CUDA, "pir_server_answer: %s", cudaGetErrorString(e));
    if (rc == PIR_OK && *(volatile uint32_t*)(h_slot + kAnswerMaxRowBytes) != 0u)
        rc = fail(PIR_ERR_CUDA, "timed out waiting for the shards' arrivals");
    if (rc == PIR_OK)
        for (uint32_t c = 0; c < W; c++) memcpy((uint8_t*)out + (size_t)c * s->slot_bytes, h_slot + (size_t)c * s->pitch, s->slot_bytes);
    lk.lock();
    s->ring_busy &= ~(1u << k);
    s->ring_cv.notify_one();
    return rc;
}
extern "C" int pir_debug_current_device(void) {  // (tests: no entry point may change the caller's current device)
    int d = -1;
    if (cudaGetDevice(&d) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    return d;
}
extern "C" void pir_debug_answer_split(uint32_t nb, uint64_t first_row, uint64_t n_rows, int num_sms, uint32_t* Dmax,
                                       uint32_t* Dmin, uint32_t* B) {
    AnswerSplit sp = choose_answer_split(nb, first_row, n_rows, num_sms);
    *Dmax = sp.D; *Dmin = sp.Dmin; *B = sp.hi_warps;
}
