#!/usr/bin/env python
"""bench.py — DPF-PIR answer throughput on B200 (BASELINE.json metric), one JSON line on rank 0.

A step = one query answered over the whole database (one pass of the hot path).  Workload at every N:
BASELINE.json configs[2], 2^24 rows x 256-byte slots (4 GiB), single query; at N > 1 the SAME
database is row-sharded over the N GPUs (strong scaling) and every shard's kernel XORs its partial
answer straight into a ring slot on rank 0's GPU (CUDA IPC mapping over NVLink; no NCCL in the data path).

  value     GB/s of database scanned per query with the database resident in HBM (device-timed)
  e2e       the same through pir_answer() with HOST buffers: key in, answer out, sync per query
  roofline  fused expand+scan kernel: N*S bytes per launch / its CUDA-event duration vs measured HBM
  cpu_baseline / --impl reference: the CPU restatement of the reference's own algorithm
            (per-row Evaluate2P + serial XOR, oracle/pir_oracle.c) on a bounded sample, all host threads
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG_ROWS = int(os.environ.get("PIRB_BENCH_LOG_ROWS", "24"))
SLOT_BYTES = int(os.environ.get("PIRB_BENCH_SLOT_BYTES", "256"))
QUERIES = int(os.environ.get("PIRB_BENCH_QUERIES", "1"))   # queries answered per step (one pir_answer_batch call)
KEYWORD = bool(int(os.environ.get("PIRB_BENCH_KEYWORD", "0")))  # configs[4]'s keyword half: 32-bit keyword queries
SEED = 2024
CPU_SAMPLE_LOG_ROWS = 20
METRIC = "dpf_pir_answer_db_scan_throughput"
UNIT = "GB/s"


def workload_name():
    return "2^%d rows x %d-byte slots (%.1f GiB DB), %s, group size 1" % (
        LOG_ROWS, SLOT_BYTES, (1 << LOG_ROWS) * SLOT_BYTES / 2**30,
        ("single keyword query (sorted distinct 32-bit keywords)" if KEYWORD else "single query") if QUERIES == 1
        else "batch of %d queries per step" % QUERIES)


def keywords_for(r0, r1):
    """Synthetic keyword column (Database.Keywords, db.go:392-394) for rows [r0, r1): sorted, distinct 32-bit values;
    row i's keyword has i in its top LOG_ROWS bits and seeded pseudo-random low bits."""
    import numpy as np
    low = 32 - LOG_ROWS
    out = np.empty(r1 - r0, dtype=np.uint64)
    for c0 in range(r0, r1, 1 << 24):   # in chunks: 2^28 keywords would otherwise need several 2 GiB temporaries
        c1 = min(r1, c0 + (1 << 24))
        i = np.arange(c0, c1, dtype=np.uint64)
        mix = (i * np.uint64(0x9E3779B97F4A7C15) + np.uint64(SEED)) >> np.uint64(40)
        out[c0 - r0:c1 - r0] = (i << np.uint64(low)) | (mix & np.uint64((1 << low) - 1))
    return out


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle is the timed thing here (the only place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup):
    import numpy as np
    import oracle
    cores = len(os.sched_getaffinity(0))
    n = 1 << LOG_ROWS
    rows = min(n, 1 << CPU_SAMPLE_LOG_ROWS, (256 << 20) // SLOT_BYTES)   # at most 256 MiB of slots
    nb = oracle.num_bits_for(n, 1, False)
    data = oracle.synth_fill(SEED, 0, rows, SLOT_BYTES)
    times = []
    for i in range(warmup + steps):
        o0, _ = oracle.gen_two_server(oracle.tape(SEED + i), nb, (i * 7919) % rows, 1)
        t0 = time.perf_counter()
        bits = oracle.expand_shared_query(o0, n, row_begin=0, n_rows=rows, nthreads=cores)  # db.go:110-174
        oracle.answer_with_bits(data, SLOT_BYTES, 1, bits)                                   # db.go:75-107
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    gbs = rows * SLOT_BYTES / (ms * 1e-3) / 1e9
    sample = "first %d of 2^%d rows (%d MiB), %d-level per-row Evaluate2P + serial XOR, per step" % (
        rows, LOG_ROWS, rows * SLOT_BYTES >> 20, nb)
    return gbs, ms, cores, sample


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    gbs, ms, cores, sample = cpu_reference_run(args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name()},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "queries_per_s": gbs * 1e9 / ((1 << LOG_ROWS) * SLOT_BYTES),
        "note": "Go toolchain absent: C restatement of the reference algorithm (oracle/), conservative (no Go allocs/goroutines)",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks sampler
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions.  The regions last milliseconds, so the
    sampler reads NVML directly from a thread (one sample per millisecond); `nvidia-smi -lms` is the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index):
        self.rows = []          # nvidia-smi csv lines (fallback)
        self.samples = []       # (sm_mhz, reasons bitmask) from NVML
        self.proc = None
        self.gpu = gpu_index
        self.nvml = None
        self.handle = None
        self.stop_flag = threading.Event()
        self.max_mhz = None

    def _nvml_open(self):
        import pynvml
        pynvml.nvmlInit()
        handle = None
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
            handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:
            handle = None
        if handle is None:
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu
            if visible:
                try:
                    idx = int(visible.split(",")[self.gpu])
                except ValueError:
                    idx = self.gpu
            handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
        self.nvml, self.handle = pynvml, handle
        self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM))

    def _poll(self):
        nv, hd = self.nvml, self.handle
        while not self.stop_flag.is_set():
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(hd, nv.NVML_CLOCK_SM))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(hd))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(hd))
                self.samples.append((mhz, mask))
            except Exception:
                pass
            time.sleep(0.001)

    def start(self):
        try:
            self._nvml_open()
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag.set()
            self.t.join(timeout=2)
            sm = sorted(m for m, _ in self.samples)
            reasons = sorted({name for _, mask in self.samples for bit, name in self.REASONS if mask & bit})
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                    "samples": len(sm), "source": "nvml, 1 ms period, during both timed regions"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 100"}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def gpu_arm(args):
    import numpy as np
    import torch
    from pir_b200 import _capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("WORLD_SIZE %d != --gpus %d" % (world, args.gpus))
    L = _capi.lib()
    if L.pir_device_count() <= local:
        raise SystemExit("bench.py: no CUDA device for local rank %d (there is no CPU fallback)" % local)
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))

    n, S = 1 << LOG_ROWS, SLOT_BYTES
    r0, r1 = rank * n // world, (rank + 1) * n // world
    h = C.c_void_p()
    _capi.check(L.pir_db_create(local, n, S, r0, r1 - r0, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, SEED))
    nb = L.pir_num_bits_for(n, 1, 1 if KEYWORD else 0)
    if KEYWORD:
        if LOG_ROWS > 32 or QUERIES != 1:
            raise SystemExit("bench.py: PIRB_BENCH_KEYWORD needs LOG_ROWS <= 32 and single queries")
        kw = np.ascontiguousarray(keywords_for(r0, r1))
        _capi.check(L.pir_db_set_keywords(h, kw.ctypes.data_as(C.POINTER(C.c_uint64)), kw.size))
        del kw

    total = args.warmup + args.steps
    rng = np.random.default_rng(SEED)
    idxs = [int(x) for x in rng.integers(0, n, size=total + 1)]
    qs = (_capi.PirQuery * (2 * (total + 1)))()
    for i, idx in enumerate(idxs):
        tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
        a = int(keywords_for(idx, idx + 1)[0]) if KEYWORD else idx
        _capi.check(L.pir_client_gen_two_server(nb, a, 1, tape, C.byref(qs[2 * i]), C.byref(qs[2 * i + 1])))
        if KEYWORD:
            qs[2 * i].is_keyword = qs[2 * i + 1].is_keyword = 1

    if QUERIES > 1 and world > 1:
        raise SystemExit("bench.py: PIRB_BENCH_QUERIES > 1 is a single-GPU workload (configs[3])")
    qbatch = None
    if QUERIES > 1:
        qbatch = (_capi.PirQuery * QUERIES)()
        scratch = _capi.PirQuery()
        for k in range(QUERIES):
            tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
            _capi.check(L.pir_client_gen_two_server(nb, int(rng.integers(0, n)), 1, tape, C.byref(qbatch[k]), C.byref(scratch)))
    stream = torch.cuda.Stream()
    d_out = torch.zeros(S * QUERIES, dtype=torch.uint8, device="cuda")
    sp = C.c_void_p(stream.cuda_stream)
    comb = None
    if world > 1:
        # one ring slot per step of either timed region plus two for the acceptance check; the slots live on
        # rank 0's GPU and every rank's kernel XORs its partial straight into them (CUDA IPC mapping, NVLink)
        from pir_b200 import sharding
        comb = sharding.open_combine_over_group(dist, local, 2 * total + 2, S, owner=0)

    def step_device(i, slot=None):
        """one query, database resident; N > 1: every rank's kernel accumulates into ring slot `slot` on rank 0"""
        if world > 1:
            comb.accumulate(h, qs[2 * i], i if slot is None else slot, sp)
        elif qbatch is not None:
            _capi.check(L.pir_answer_batch_device(h, qbatch, QUERIES, C.c_void_p(d_out.data_ptr()), sp))
        else:
            _capi.check(L.pir_answer_batch_device(h, C.byref(qs[2 * i]), 1, C.c_void_p(d_out.data_ptr()), sp))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- acceptance check (outside the timed region): both shares recover the queried slot ----
    i = total
    host = np.zeros(S, dtype=np.uint8)
    if world == 1:
        acc = torch.zeros(S, dtype=torch.uint8, device="cuda")
        for share in (0, 1):
            _capi.check(L.pir_answer_batch_device(h, C.byref(qs[2 * i + share]), 1, C.c_void_p(d_out.data_ptr()), sp))
            stream.synchronize()
            acc ^= d_out[:S]
        recovered = acc.cpu().numpy()
    else:
        for share in (0, 1):
            comb.accumulate(h, qs[2 * i + share], 2 * total + share, sp)
        barrier()
        recovered = None
        if rank == 0:
            recovered = comb.wait_read(2 * total, world, S) ^ comb.wait_read(2 * total + 1, world, S)
    box = [None]
    owner_rank = next(r for r in range(world) if r * n // world <= idxs[i] < (r + 1) * n // world)
    if rank == owner_rank:
        _capi.check(L.pir_db_download(h, idxs[i], host.ctypes.data_as(C.c_void_p), 1))
        box[0] = host.tobytes()
    if world > 1:
        dist.broadcast_object_list(box, src=owner_rank)
    if rank == 0 and not os.environ.get("PIRB_BENCH_NOCHECK"):
        if not np.array_equal(recovered, np.frombuffer(box[0], dtype=np.uint8)):
            raise SystemExit("bench.py: recovered slot differs from the database slot — refusing to time a wrong path")

    # ---- warm-up ----
    for i in range(args.warmup):
        step_device(i)
    barrier()

    # ---- device-resident timed region ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    _capi.check(L.pir_db_set_kernel_timing(h, 1))
    launches0 = L.pir_kernel_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for i in range(args.warmup, total):
        step_device(i)
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1)
    launches = L.pir_kernel_launch_count() - launches0
    kms, kcount = C.c_double(), C.c_uint64()
    _capi.check(L.pir_db_get_kernel_timing(h, C.byref(kms), C.byref(kcount)))
    _capi.check(L.pir_db_set_kernel_timing(h, 0))

    # ---- end-to-end timed region: host key in, host answer out, every step ----
    out_host = np.zeros(S * QUERIES, dtype=np.uint8)
    if world == 1:
        for i in range(min(3, args.warmup)):
            if qbatch is not None:
                _capi.check(L.pir_answer_batch(h, qbatch, QUERIES, out_host.ctypes.data_as(C.c_void_p), S * QUERIES))
            else:
                _capi.check(L.pir_answer(h, C.byref(qs[2 * i]), out_host.ctypes.data_as(C.c_void_p), S))
    barrier()
    t0 = time.perf_counter()
    for i in range(args.warmup, total):
        if qbatch is not None:
            tq = time.perf_counter()
            _capi.check(L.pir_answer_batch(h, qbatch, QUERIES, out_host.ctypes.data_as(C.c_void_p), S * QUERIES))
            if os.environ.get("PIRB_BENCH_VERBOSE"):
                print("e2e batch call %.2f ms" % ((time.perf_counter() - tq) * 1e3), file=sys.stderr)
        elif world == 1:
            _capi.check(L.pir_answer(h, C.byref(qs[2 * i]), out_host.ctypes.data_as(C.c_void_p), S))
        else:
            # every rank: host key -> its shard's kernel -> XOR into ring slot on rank 0;
            # rank 0: device-side wait for all N arrivals, then the combined answer comes back to the host
            comb.accumulate(h, qs[2 * i], total + i, None)
            if rank == 0:
                out_host = comb.wait_read(total + i, world, S)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms, kms.value / max(1, kcount.value)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms, k_avg_ms = [float(x) for x in t.cpu()]
    else:
        k_avg_ms = kms.value / max(1, kcount.value)
    if k_avg_ms <= 0.0:
        # the two-phase batch path does not time single launches: use the step time per query
        k_avg_ms = dev_ms / args.steps / QUERIES

    if rank == 0:
        db_bytes = n * S * QUERIES   # a batch makes one pass over the database per query
        ms_step = dev_ms / args.steps
        value = db_bytes / (ms_step * 1e-3) / 1e9
        e2e_step = e2e_ms / args.steps
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        shard_bytes = (r1 - r0) * S
        achieved = shard_bytes / (k_avg_ms * 1e-3) / 1e9
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["answer_kernel"]
            if world == tr["n_gpus"] and tr["workload"] in workload_name():
                traffic = tr["dram_bytes_read"] + tr["dram_bytes_write"]
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(), "sharding": "rows/%d" % world,
                       "l2": "per-GPU shard (%d MiB) exceeds the 126 MB L2; every step scans all of it" % (shard_bytes >> 20),
                       "combine": "none" if world == 1 else "every shard's kernel XORs its %d-byte partial into rank 0's ring slot "
                                  "(red.global.xor over a CUDA IPC mapping / NVLink); no NCCL in the data path" % S,
                       "queries": "distinct seeded key per step, share 0; FinalCW even / odd in the timed steps: %d / %d "
                                  "(a coin flip per key; the kernel has a code path per parity)" % (
                                      sum(1 for i in range(args.warmup, total) if not qs[2 * i].final_cw & 1),
                                      sum(1 for i in range(args.warmup, total) if qs[2 * i].final_cw & 1))},
            "queries_per_s": 1e3 * QUERIES / ms_step,
            "e2e": {"value": db_bytes / (e2e_step * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_step,
                    "h2d_bytes_per_step": C.sizeof(_capi.PirQuery) * QUERIES, "d2h_bytes_per_step": S * QUERIES,
                    "api": ("pir_answer (C ABI, host key in / host answer out, stream sync per query)" if QUERIES == 1 else
                            "pir_answer_batch (C ABI, host keys in / host answers out, one sync per batch)") if world == 1 else
                           "pir_answer_accumulate_device on every rank + pir_combine_wait_read on rank 0 (host key in, combined host answer out)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "kernel": "keyword_walk_kernel + scan_bytes_kernel" if KEYWORD else "answer_kernel" if QUERIES == 1 else "answer_kernel (bit vectors) + batch_scan_kernel, per query", "kernel_ms": k_avg_ms,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650",
                         "algorithmic_bytes_per_launch": shard_bytes,
                         "aes_blocks_per_s": None if KEYWORD else 3.0 * (r1 - r0) / (k_avg_ms * 1e-3)},
            "clocks": clocks,
        }
        if world == 1:
            gbs, ms, cores, sample = cpu_reference_run(3, 1)
            line["cpu_baseline"] = {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if comb is not None:
        barrier()
        if rank != 0:
            comb.close()   # peers unmap before the owner frees
        dist.barrier()
        if rank == 0:
            comb.close()
    L.pir_db_destroy(h)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
