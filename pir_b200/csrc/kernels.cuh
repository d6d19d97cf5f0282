// kernels.cuh — launch-side declarations shared by kernels.cu (device code) and capi.cu (C ABI).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "dpf_device.cuh"

namespace pirb {

constexpr int kFusedThreads = 512;  // threads per CTA of the fused kernel (one CTA per SM)
constexpr int kFusedLogThreads = 9;
constexpr int kMaxDfsDepth = 7;     // per-thread DFS depth D: a thread owns 2^(D+1) rows
constexpr int kTopThreads = 1024;

// Geometry of one fused launch over a shard (all row numbers are GLOBAL row indices of the DPF
// domain; the shard's first row is stored at `db`).
struct FusedPlan {
    const uint8_t* db;        // shard base (row `first_row` lives here)
    uint64_t row_bytes;       // W * slot_pitch, multiple of 16
    uint64_t first_row;       // first global row of the shard
    uint64_t n_rows;          // rows in the shard
    uint32_t D;               // per-thread DFS depth
    uint32_t B;               // CTA-level BFS levels: 2^B threads get a subtree each
    uint32_t Fc;              // tree level of the CTA tasks = num_bits-1-D-B
    uint32_t pad;
    uint64_t task_first;      // first CTA task (node index at level Fc) touching the shard
    uint64_t task_count;
    const uint4* frontier_s;  // task_count seeds at level Fc (from the top kernel)
    const uint8_t* frontier_t;
    uint32_t* out;            // SCAN mode: row_bytes/4 words, zeroed before launch (atomic XOR)
    uint8_t* bits_out;        // BITS mode: one byte per shard row
};

// ---- look-ahead of the next query's tree top (answer_kernel.cu "look-ahead") ---------------------------------------
// The top of a task is a chain of ~20 dependent node expansions that nothing in the SAME query can hide.  Under load
// the next query of the handle is already enqueued while this one runs, so the host leaves its key material in a
// host-mapped ring (LaMail, written at enqueue time), the walker warp of launch j computes the top of launch j+1's
// task for the same CTA and leaves the nodes in global memory (LaTop), and launch j+1 starts from them.  Purely
// opportunistic: every entry is validated by its sequence number (mail: seqlock; tops: release/acquire), and a
// launch that finds nothing walks its own top as before.
constexpr uint32_t kLaMailRing = 64;   // launches the host may run ahead of the GPU before look-ahead stops finding its mail
constexpr uint32_t kLaTopRing = 4;     // launch j+1's entry is written during launch j and read at the start of j+1
constexpr uint32_t kLaMaxLevels = 32;  // correction words kept per mail entry (levels above the thread roots)
constexpr uint32_t kLaHand = 128;      // nodes handed over per CTA (answer_kernel.cu kHand)
struct LaConst {                       // what a top walk needs of QueryConst; same member names
    uint32_t rk[3][44];
    uint32_t cw[kLaMaxLevels][4];
    uint64_t cw_tl, cw_tr;
    uint32_t s_init[4];
    uint32_t t_init, pad;
};
struct LaMail {
    uint64_t tag;    // sequence number of the launch this entry belongs to; 0 while it is being written
    uint64_t geom;   // hash of shard geometry and work split: the top is only valid for an identical launch shape
    uint64_t claim;  // device copy only: highest sequence number some CTA has set out to fetch from the host ring
    LaConst c;
};
struct LaTop {       // one per (ring entry, CTA)
    uint64_t ready;  // sequence number of the launch these nodes are for
    uint32_t n, lh;  // nodes, tree level
    uint4 seed[kLaHand];
    uint8_t t[kLaHand];
};

// Geometry of one launch of the warp-specialised answer kernel (answer_kernel.cu).  Same meaning as
// FusedPlan, but the kernel walks the top of the tree itself (no frontier buffers).
struct AnswerPlan {
    const uint8_t* db;
    uint64_t row_bytes;   // power of two, 16..512
    uint64_t first_row;
    uint64_t n_rows;
    // work split (answer_kernel.cu "work split"):
    uint32_t D;           // uniform: every task has the shape (D, hi_warps): threads of the first hi_warps tree
    uint32_t hi_warps;    //   warps own 2^(D+1) rows each, the others 2^D; hi_warps = 4, 8, 12 or 16 (16 if D = 0).
    uint32_t Dmin;        // binary (hi_warps == 0): chunks of 2^(10+Dmin) rows covered by aligned tasks of depth Dmin..D
    uint32_t pad;         // 0 (PIRB_DEV build: experiment flags, dev_hooks.cuh)
    uint32_t m_task;      // uniform split: level-L nodes (2^(D+1) rows) per task if fewer than the shape's own (0 = the shape's):
    uint32_t pad2;        //   tasks trimmed so that all SMs get the same number of rows
    uint32_t slog_cap;    // log2(rows per thread per slice) at most: 3, or 2 when tasks are so shallow that the last
    uint32_t top_warp;    //   slice (whose rows are scanned after the tree warps are done) would be a large part of a CTA's work
                          // top_warp != 0: an extra warp walks the top of task k+1 during the depth-first phase of task k
    uint32_t* out;        // row_bytes/4 words, XOR-accumulated (atomics); the caller zeroes
    uint32_t* bits_out;   // BITS mode (non-null): packed selection bits, bit i = row first_row + i, OR-accumulated
                          // into a zeroed buffer of ceil(n_rows/32)+1 words; no rows are read
#ifdef PIRB_DEV
    unsigned long long* dbg;  // per-CTA phase cycle counters (PIRB_PHASE_DEBUG), else null; DEV build only
#endif
    // multi-GPU combine (optional): the last CTA to finish bumps *arrivals with system scope, after every CTA's
    // partial has been XORed into `out` (which may be peer memory); cta_done is a zeroed counter in local memory
    uint32_t* arrivals;
    uint32_t* cta_done;
    // single-GPU call (optional): the launch's last CTA moves the finished answer from `out` to this buffer (device
    // memory or a mapped pinned host buffer) and leaves `out` zeroed for the next query -- no memset before and no
    // copy after the kernel.  Needs cta_done.  With `arrivals` set as well, final_out is a multi-GPU combine slot: the last CTA
    // XORs the shard's sum into it (system scope) and then bumps *arrivals.
    uint32_t* final_out;
    // look-ahead (optional, all null / 0 = off): this launch's sequence number on its handle, the hash of its shape,
    // the host-mapped mail ring [kLaMailRing] and the tops ring [kLaTopRing][gridDim.x] in device memory
    uint64_t la_seq, la_geom;
    const LaMail* la_mail;
    LaMail* la_dmail;     // [kLaTopRing] in device memory: CTA 0 republishes the next launch's mail here for the other CTAs
    LaTop* la_tops;
#ifdef PIRB_TEX_ROWS
    // experiment: rows fetched through the texture path; a linear texture of uint4 texels over `db`
    unsigned long long tex;
#endif
};

// The in-kernel phase counters (clock64 / globaltimer) and the experiment flags exist in the PIRB_DEV build only
// (make dev); in the shipped build these are compile-time constants and the code behind them is dropped.
#ifdef PIRB_DEV
#define PIRB_AP_DBG(ap) ((ap).dbg)
#define PIRB_AP_PAD(ap) ((ap).pad)
#else
#define PIRB_AP_DBG(ap) (static_cast<unsigned long long*>(nullptr))
#define PIRB_AP_PAD(ap) 0u
#endif

// Batched scan (batch_scan.cu): rows [0, n_rows) of a shard against n_queries packed bit vectors.
struct BatchScanPlan {
    const uint8_t* db;      // shard base
    uint64_t row_bytes;     // multiple of 16, <= 32 KB
    uint64_t n_rows;
    const uint32_t* bits;   // [n_queries][wpq] packed selection bits (bit i of a vector = shard row i)
    uint64_t wpq;           // words per query vector (>= ceil(n_rows / 32))
    uint32_t n_queries;     // <= queries_per_launch of batch_scan_geometry
    uint32_t tpq;           // threads per query (power of two)
    uint32_t seg_chunks;    // 16-byte chunks per thread (<= 8)
    uint32_t stage_rows;    // rows per shared-memory stage (power of two <= 32)
    uint32_t rsplit;        // row classes (power of two <= 32): n_queries * tpq * rsplit <= 512
    uint32_t* out;          // [n_queries][row_bytes / 4] words, XOR-accumulated; the caller zeroes
};

struct TopPlan {
    uint32_t Fc;          // produce nodes of level Fc
    uint32_t pad;
    uint64_t task_first;  // node index range wanted at level Fc
    uint64_t task_count;
    uint4* buf_s[2];      // ping-pong level buffers, each >= task_count + 2 entries
    uint8_t* buf_t[2];
    // the result is left in buf_s[Fc & 1], entry j = node task_first + j
};

// row layout kinds handled by the fused kernel: chunks of 16 B per row
//   c = 1,2,4,8,16,32  -> one accumulator per lane
//   c = 64,128,256     -> c/32 accumulators per lane
bool fused_supports_row_bytes(uint64_t row_bytes);

cudaError_t launch_top(const QueryConst& qc, const TopPlan& tp, const uint32_t* d_t0, uint32_t smem_base,
                       cudaStream_t st);
cudaError_t launch_fused(const QueryConst& qc, const FusedPlan& fp, bool bits_mode, const uint32_t* d_t0,
                         uint32_t smem_base, int num_sms, cudaStream_t st);
// per-point evaluation (Evaluate2P): values (int64) and/or selection bits (bytes) for n points.
// xs == nullptr means x = x_first + i.
cudaError_t launch_eval_points(const QueryConst& qc, const uint64_t* xs, uint64_t x_first, uint64_t n,
                               int64_t* values, uint8_t* bits, const uint32_t* d_t0, uint32_t smem_base,
                               cudaStream_t st);
// the same with prefix sharing between consecutive points (keyword_kernel.cu): the keyword path
cudaError_t configure_keyword_kernel();
// db != nullptr: the row XOR is fused into the walk (rows of a multiple of 16 up to 512 bytes, keyword_fused_supports)
bool keyword_fused_supports(uint64_t row_bytes);
cudaError_t launch_keyword_walk(const QueryConst& qc, const uint64_t* xs, uint64_t x_first, uint64_t n, int64_t* values,
                                uint8_t* bits, const uint32_t* d_t0, uint32_t smem_base, int num_sms, cudaStream_t st,
                                const uint8_t* db = nullptr, uint64_t row_bytes = 0, uint32_t* out = nullptr);
// masked XOR of rows from a byte-per-row bit vector (any row_bytes multiple of 16)
cudaError_t launch_scan_bytes(const uint8_t* db, uint64_t row_bytes, uint64_t n_rows, const uint8_t* bits,
                              uint32_t* out, int num_sms, cudaStream_t st);
cudaError_t launch_synth_fill(uint8_t* db, uint64_t slot_begin, uint64_t n_slots, uint32_t slot_bytes,
                              uint32_t pitch, uint64_t seed, cudaStream_t st);
cudaError_t probe_smem_base(uint32_t* out_base);
cudaError_t configure_kernels();

// warp-specialised fused expand+scan kernel (answer_kernel.cu)
bool answer_kernel_supports(uint64_t row_bytes, uint32_t D);
cudaError_t configure_answer_kernel();
uint64_t answer_tasks(uint64_t first_row, uint64_t n_rows, uint32_t D, uint32_t hi_warps, uint32_t m_task = 0);  // tasks of a uniform split
uint32_t answer_task_nodes(uint32_t hi_warps, uint32_t m_task);  // level-L nodes per task of a uniform split
uint64_t answer_chunks(uint64_t first_row, uint64_t n_rows, uint32_t Dmin);                  // chunks of a binary split
// overlap: programmatic dependent launch (the kernel's CTAs may start while the previous kernel in the stream drains;
// all its global writes wait for that kernel's completion)
cudaError_t launch_answer(const QueryConst& qc, const AnswerPlan& ap, const uint32_t* d_t0, uint32_t smem_base,
                          int num_sms, cudaStream_t st, bool overlap = false);

// packed bits (bit i = row i) -> one 0/1 byte per row; `bytes` must be 4-byte aligned
cudaError_t launch_unpack_bits(const uint32_t* packed, uint8_t* bytes, uint64_t n_rows, cudaStream_t st);
// multi-GPU combine signalling: bump / wait for a slot's arrival counter (system scope)
cudaError_t launch_signal_arrival(uint32_t* arrivals, cudaStream_t st);
cudaError_t launch_wait_arrivals(const uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles,
                                 uint32_t* timed_out, cudaStream_t st);
// wait, move the combined answer to a mapped pinned host buffer, zero slot and counter: one launch
cudaError_t launch_wait_read_reset(uint32_t* arrivals, uint32_t expected, uint64_t timeout_cycles, uint32_t* slot, uint32_t* h_out,
                                   uint32_t* h_flag, uint32_t words, cudaStream_t st);

bool batch_scan_geometry(uint64_t row_bytes, uint32_t* tpq, uint32_t* seg_chunks, uint32_t* stage_rows,
                         uint32_t* queries_per_launch);
cudaError_t configure_batch_scan();
cudaError_t launch_batch_scan(const BatchScanPlan& bp, int num_sms, cudaStream_t st);

uint64_t launch_count();
void count_launch();

}  // namespace pirb
