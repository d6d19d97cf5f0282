// dev_hooks.cuh — development instrumentation, compiled ONLY into the -DPIRB_DEV build (make dev ->
// pir_b200/lib/libpir_b200_dev.so, used by tools/).  The shipped library (make all) contains none of this: it
// reads no environment variable, its kernels carry no clock64()/globaltimer counters and no switch that skips
// work.  In the DEV build the tuning options of pir_db_set_option can also come from PIRB_* environment
// variables (re-read at every query, so the sweeps in tools/ can flip them between calls), and three
// experiment switches exist that DO change results or skip work: PIRB_DEBUG_FLAGS (bit 0: no row loads, bit 1:
// force the odd-FinalCW code path), PIRB_BATCH_SKIP_SCAN (phase 1 only), PIRB_PHASE_DEBUG (per-CTA counters).
#pragma once
#ifndef PIRB_DEV
#error "dev_hooks.cuh is for the PIRB_DEV build only"
#endif
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace pirb_dev {

inline const char* dev_env(const char* name) { return getenv(name); }

template <class T>
inline void dev_apply_env(T& t) {
    auto num = [](const char* n, int dflt) { const char* e = getenv(n); return e ? atoi(e) : dflt; };
    t.dfs_depth = num("PIRB_DFS_DEPTH", -1);
    t.hi_warps = num("PIRB_HI_WARPS", -1);
    t.dmin = num("PIRB_DMIN", -1);
    t.top_warp = num("PIRB_TOP_WARP", -1);
    t.slice_log = num("PIRB_SLICE_LOG", 0);
    const char* bm = getenv("PIRB_BATCH_MODE");
    t.batch_mode = !bm ? 0 : !strcmp(bm, "fused") ? 1 : !strcmp(bm, "bits") ? 2 : 0;
    t.batch_one_stream = getenv("PIRB_BATCH_ONE_STREAM") != nullptr;
    t.batch_latency_split = getenv("PIRB_BATCH_LATENCY_SPLIT") != nullptr;
    t.kernel_v1 = getenv("PIRB_KERNEL") && !strcmp(getenv("PIRB_KERNEL"), "v1");
    t.expand_v1 = getenv("PIRB_EXPAND_V1") != nullptr;
    t.keyword_v1 = getenv("PIRB_KEYWORD_V1") != nullptr;
    t.force_unfused = getenv("PIRB_FORCE_UNFUSED") != nullptr;
    t.signal_kernel = getenv("PIRB_SIGNAL_KERNEL") != nullptr;
    t.host_copy_path = getenv("PIRB_HOST_COPY_PATH") != nullptr;
    t.serial_launch = getenv("PIRB_SERIAL_LAUNCH") != nullptr;
}

inline bool dev_phase_debug() { return getenv("PIRB_PHASE_DEBUG") != nullptr; }
inline bool dev_skip_scan() { return getenv("PIRB_BATCH_SKIP_SCAN") != nullptr; }

#define DEV_CU(call)                                                                   \
    do {                                                                               \
        cudaError_t e_ = (call);                                                       \
        if (e_ != cudaSuccess) {                                                       \
            fprintf(stderr, "[dev] %s: %s\n", #call, cudaGetErrorString(e_));          \
            return -3;                                                                 \
        }                                                                              \
    } while (0)

// per-CTA phase counters of one answer launch (answer_kernel.cu, ap.dbg): allocate before, print after
struct DevPhase {
    unsigned long long* d_dbg = nullptr;
    int begin(pirb::AnswerPlan* ap, cudaStream_t st) {
        if (const char* dbg = getenv("PIRB_DEBUG_FLAGS")) ap->pad = (uint32_t)atoi(dbg);  // timing experiments only
        if (!dev_phase_debug()) return 0;
        DEV_CU(cudaMalloc(&d_dbg, 8 * 15 * 1024));
        DEV_CU(cudaMemsetAsync(d_dbg, 0, 8 * 15 * 1024, st));
        ap->dbg = d_dbg;
        return 0;
    }
    int report(cudaStream_t st) {
        if (!d_dbg) return 0;
  // development aid: average per-CTA cycles per phase of the tree warps, to stderr
            std::vector<unsigned long long> h(15 * 1024);
            DEV_CU(cudaStreamSynchronize(st));
            DEV_CU(cudaMemcpy(h.data(), d_dbg, 8 * 15 * 1024, cudaMemcpyDeviceToHost));
            cudaFree(d_dbg);
            double sum[8] = {0}, mx = 0;
            int n = 0;
            for (int b = 0; b < 1024; b++) {
                if (!h[8 * b + 5]) continue;
                n++;
                for (int k = 0; k < 8; k++) sum[k] += (double)h[8 * b + k];
                if ((double)(h[8 * b + 5] + h[8 * b + 7]) > mx) mx = (double)(h[8 * b + 5] + h[8 * b + 7]);
            }
            unsigned long long g0 = ~0ull, g1 = 0, gs_last = 0, ge_first = ~0ull;
            for (int b = 0; b < 1024; b++)
                if (h[8 * b + 5]) {
                    if (h[12288 + b] < g0) g0 = h[12288 + b];
                    if (h[12288 + b] > gs_last) gs_last = h[12288 + b];
                    if (h[13312 + b] > g1) g1 = h[13312 + b];
                    if (h[13312 + b] < ge_first) ge_first = h[13312 + b];
                }
            if (dev_env("PIRB_PHASE_DEBUG")[0] == '3')
                for (int l = 1; l < 64; l++)
                    if (h[14336 + l] && h[14336 + l - 1]) fprintf(stderr, "[top cta0] level %d took %llu cycles\n", l, h[14336 + l] - h[14336 + l - 1]);
            if (n) fprintf(stderr, "[phase] globaltimer: first CTA start to last CTA start %llu ns, first CTA end at %llu ns, last CTA end at %llu ns\n",
                           gs_last - g0, ge_first - g0, g1 - g0);
            if (dev_env("PIRB_PHASE_DEBUG")[0] == '2')
                for (int b = 0; b < 1024; b++)
                    if (h[8 * b + 5])
                        fprintf(stderr, "[cta %d] tasks %llu prologue %llu top %llu bfs %llu dfs %llu tail %llu total %llu start_ns %llu end_ns %llu\n", b, h[8 * b + 6],
                                h[8 * b + 7], h[8 * b + 0], h[8 * b + 1], h[8 * b + 3], h[8 * b + 4], h[8 * b + 5] + h[8 * b + 7],
                                h[12288 + b] - g0, h[13312 + b] - g0);
            double spread = 0;
            for (int b = 0; b < 1024; b++)
                if (h[8 * b + 5]) spread += (double)h[8192 + 2 * b] - (double)((1ull << 40) - h[8192 + 2 * b + 1]);
            double busy = 0, waitc = 0;
            for (int b = 0; b < 1024; b++)
                if (h[8 * b + 5]) {
                    busy += (double)(h[11264 + b] >> 32);
                    waitc += (double)(h[11264 + b] & 0xffffffffull);
                }
            if (n) fprintf(stderr, "[phase] first scan warp: %.0f cycles scanning, %.0f waiting for slices\n", busy / n, waitc / n);
            double backlog = 0;
            for (int b = 0; b < 1024; b++)
                if (h[8 * b + 5]) backlog += (double)h[10240 + b] - 1.0;
            if (n) fprintf(stderr, "[phase] scan backlog when a tree warp finishes (max over its warps): %.1f units of 32 words\n", backlog / n);
            if (n) fprintf(stderr, "[phase] first-to-last tree warp done within a CTA: %.0f cycles (mean)\n", spread / n);
            if (n)
                fprintf(stderr, "[phase] ctas %d tasks %.2f | prologue %.0f top %.0f bfs %.0f dfs %.0f (of which ring-full wait %.0f) tail %.0f | tree total %.0f cycles, slowest CTA %.0f\n",
                        n, sum[6] / n, sum[7] / n, sum[0] / n, sum[1] / n, sum[3] / n, sum[2] / n, sum[4] / n, sum[5] / n, mx);
        
        return 0;
    }
};

}  // namespace pirb_dev
