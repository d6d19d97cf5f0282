#!/usr/bin/env python
"""bench.py — DPF-PIR answer throughput on B200 (BASELINE.json metric), one JSON line on rank 0.

A step = one query answered over the whole database (one pass of the hot path).  Workload at every N:
BASELINE.json configs[2], 2^24 rows x 256-byte slots (4 GiB), single query; at N > 1 the SAME database is
row-sharded over the N GPUs (strong scaling) and every shard's kernel XORs its partial answer straight into a
ring slot on rank 0's GPU (CUDA IPC mapping over NVLink; no NCCL in the data path).

  value     GB/s of database scanned per query, database resident in HBM, device-timed: the median of 3 regions of exactly
            `steps` steps (all three listed in config.value_regions_ms); a region = `steps` single-query
            launches back to back on one stream per GPU (consecutive launches overlap as programmatic dependent
            launches; at N > 1 rank 0's event pair also covers the wait for the last query's N arrivals)
  e2e       the same through the C ABI with HOST buffers and queries in flight (like `value`, which is back-to-back launches):
            N = 1: pir_answer from 8 host threads on one handle; N > 1: the loop below with 4 queries in flight.
  e2e_one_at_a_time  key in host memory -> answer in host memory, the next query only after the answer.  N = 1: pir_answer.  N > 1: the key lives in rank 0's host memory only and is fanned out to the
            other processes through a shared-memory mailbox; every rank launches its shard; rank 0 waits for the N
            arrivals, reads the combined answer and re-zeroes the ring slot (ring of 4) -- all inside the timed loop.
            `e2e_single_process` times pir_server_answer (one process drives all N GPUs: the Go server's path).
  roofline  dominant kernel vs measured HBM bandwidth on ALGORITHMIC bytes (rows x row_bytes per launch), plus the
            bound that binds: `lookup` = AES blocks/s against the AES-only rate of the same shard measured in this
            run (BITS-mode launches: same tree expansion, no row read), and `dram` = physical bytes (selected rows
            only: popcount of the step's selection bits x row_bytes) per second.
  parity    every timed answer (device-timed and e2e regions, both servers' shares alternating) is compared with the
            oracle's bytes after the timed regions: bits by oracle.expand_tree, rows by oracle.answers_over_synth.
  other_configs  BASELINE configs[1], [3], [4] (index and keyword) on the shipped kernels, each verified.
  cpu_baseline / --impl reference: the CPU restatement of the reference's own algorithm (per-row Evaluate2P + serial
            XOR, oracle/pir_oracle.c) on a bounded sample, all host threads.
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
OTHER = bool(int(os.environ.get("PIRB_BENCH_OTHER", "1")))      # append other_configs to the line
SEED = 2024
CPU_SAMPLE_LOG_ROWS = 20
METRIC = "dpf_pir_answer_db_scan_throughput"
UNIT = "GB/s"
CONC = 8            # host threads of the concurrent end-to-end region (N = 1)
CONC_PASSES = 4     # passes of those threads over the timed keys
PIPE_DEPTH = 4      # queries in flight in the pipelined end-to-end region (N > 1)
PIPE_RING = 8       # its combine slots (a slot is re-used 8 queries later, after its answer has been read and it was re-zeroed)
REGIONS = 3         # timed regions of exactly `steps` steps each; the median is reported, all are listed
RING = 4            # combine slots of the e2e loop at N > 1 (reset and re-used inside the timed loop)


def workload_name(log_rows=None, slot_bytes=None, queries=1, keyword=False):
    log_rows = LOG_ROWS if log_rows is None else log_rows
    slot_bytes = SLOT_BYTES if slot_bytes is None else slot_bytes
    return "2^%d rows x %d-byte slots (%.1f GiB DB), %s, group size 1" % (
        log_rows, slot_bytes, (1 << log_rows) * slot_bytes / 2**30,
        ("single keyword query (sorted distinct 32-bit keywords)" if keyword else "single query") if queries == 1
        else "batch of %d queries per step" % queries)


def keywords_for(r0, r1, log_rows):
    """Synthetic keyword column (Database.Keywords, db.go:392-394) for rows [r0, r1): sorted, distinct 32-bit values;
    row i's keyword has i in its top log_rows bits and seeded pseudo-random low bits."""
    import numpy as np
    low = 32 - log_rows
    out = np.empty(r1 - r0, dtype=np.uint64)
    for c0 in range(r0, r1, 1 << 24):   # in chunks: 2^28 keywords would otherwise need several 2 GiB temporaries
        c1 = min(r1, c0 + (1 << 24))
        i = np.arange(c0, c1, dtype=np.uint64)
        mix = (i * np.uint64(0x9E3779B97F4A7C15) + np.uint64(SEED)) >> np.uint64(40)
        out[c0 - r0:c1 - r0] = (i << np.uint64(low)) | (mix & np.uint64((1 << low) - 1))
    return out


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle is the timed thing here (the only place bench.py executes oracle/ as the measured path)
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup):
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
    sample = ("first %d of 2^%d rows (%d MiB), %d-level per-row Evaluate2P over %d threads (rows split statically, not "
              "the reference's goroutine-per-row fan-out of db.go:149-170) + serial XOR, per step" % (
                  rows, LOG_ROWS, rows * SLOT_BYTES >> 20, nb, cores))
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
        "note": "Go toolchain absent: C restatement of the reference algorithm (oracle/), static row split; conservative "
                "(no per-call Go allocations, no goroutine per row)",
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
class Env:
    """process-group plumbing shared by the headline workload and the other configs"""

    def __init__(self, args):
        import numpy as np
        import torch
        from pir_b200 import _capi, sharding
        self.np, self.torch, self.capi, self.sharding = np, torch, _capi, sharding
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit("WORLD_SIZE %d != --gpus %d" % (self.world, args.gpus))
        self.L = _capi.lib()
        if self.L.pir_device_count() <= self.local:
            raise SystemExit("bench.py: no CUDA device for local rank %d (there is no CPU fallback)" % self.local)
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group(backend="nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
            self.cpu_group = dist.new_group(backend="gloo")   # host-only barrier: an NCCL barrier is a kernel spinning on the GPU
        self.stream = torch.cuda.Stream()
        self.sp = C.c_void_p(self.stream.cuda_stream)
        self.cores = len(os.sched_getaffinity(0))

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def cpu_barrier(self):
        """wait without touching the GPU (the other ranks' GPUs must be idle while rank 0 drives them itself)"""
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def max_over_ranks(self, vals):
        if self.world == 1:
            return [float(v) for v in vals]
        t = self.torch.tensor(vals, dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def shard(self, n):
        return self.rank * n // self.world, (self.rank + 1) * n // self.world

    def make_handle(self, n, S, seed):
        r0, r1 = self.shard(n)
        h = C.c_void_p()
        self.capi.check(self.L.pir_db_create(self.local, n, S, r0, r1 - r0, C.byref(h)))
        self.capi.check(self.L.pir_db_fill_synthetic(h, seed))
        return h, r0, r1

    def gen_keys(self, rng, nb, targets, keyword=False):
        """both servers' keys for every target: PirQuery array [2 * len(targets)] (share s of key k at 2k + s)"""
        np, capi = self.np, self.capi
        qs = (capi.PirQuery * (2 * len(targets)))()
        for k, a in enumerate(targets):
            tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
            capi.check(self.L.pir_client_gen_two_server(nb, int(a), 1, tape, C.byref(qs[2 * k]), C.byref(qs[2 * k + 1])))
            if keyword:
                qs[2 * k].is_keyword = qs[2 * k + 1].is_keyword = 1
        return qs


def oracle_answers(env, n, S, seed, queries):
    """The oracle's answer bytes for product keys over the synthetic database: bits by tree expansion of the restated
    Evaluate2P (db.go:128-147), rows XORed as db.go:91-104 does, generated on the fly.  Returns (answers, mean popcount)."""
    import oracle
    np = env.np
    out, pops = [], []
    for g0 in range(0, len(queries), 8):
        grp = [oracle.query_from_bytes(q.to_bytes()) for q in queries[g0:g0 + 8]]
        bits = [oracle.expand_tree(o, n, nthreads=env.cores) for o in grp]
        pops += [int(np.count_nonzero(b)) for b in bits]
        out.append(oracle.answers_over_synth(seed, S, 1, bits, nthreads=env.cores))
    return np.concatenate(out), float(np.mean(pops))


class Mailbox:
    """Query fan-out for the one-process-per-GPU layout: the key arrives in rank 0's host memory only; rank 0 writes it
    (1256 bytes) into slot seq % SLOTS of a POSIX shared-memory ring and then bumps a sequence number; the other ranks
    poll the sequence number.  A slot is re-used SLOTS queries later, which the caller's flow control makes safe."""
    SLOTS = 8

    def __init__(self, env, qbytes):
        from multiprocessing import shared_memory
        np = env.np
        name = "pirb_bench_%s" % os.environ.get("MASTER_PORT", "0")
        self.env, self.qbytes = env, qbytes
        self.pitch = (qbytes + 63) // 64 * 64
        size = self.SLOTS * self.pitch + 64        # the sequence number sits in its own cache line behind the keys
        if env.rank == 0:
            try:
                old = shared_memory.SharedMemory(name=name)
                old.close()
                old.unlink()
            except FileNotFoundError:
                pass
            self.shm = shared_memory.SharedMemory(name=name, create=True, size=size)
            self.shm.buf[:size] = bytes(size)
        env.barrier()
        if env.rank != 0:
            self.shm = shared_memory.SharedMemory(name=name)
        self.payload = np.ndarray((self.SLOTS, self.pitch), dtype=np.uint8, buffer=self.shm.buf, offset=0)
        self.seq = np.ndarray((1,), dtype=np.uint64, buffer=self.shm.buf, offset=self.SLOTS * self.pitch)
        env.barrier()

    def publish(self, q, seq):
        self.payload[seq % self.SLOTS, :self.qbytes] = self.env.np.frombuffer(q.to_bytes(), dtype=self.env.np.uint8)
        self.seq[0] = seq

    def receive(self, seq, timeout_s=30.0):
        t0 = time.perf_counter()
        while int(self.seq[0]) < seq:
            if time.perf_counter() - t0 > timeout_s:
                raise SystemExit("bench.py: rank %d timed out waiting for query %d" % (self.env.rank, seq))
        return self.env.capi.PirQuery.from_bytes(self.payload[seq % self.SLOTS, :self.qbytes].tobytes())

    def close(self):
        del self.payload, self.seq
        self.shm.close()
        self.env.barrier()
        if self.env.rank == 0:
            self.shm.unlink()


def headline(env, args):
    np, torch, capi, L = env.np, env.torch, env.capi, env.L
    world, rank = env.world, env.rank
    n, S = 1 << LOG_ROWS, SLOT_BYTES
    h, r0, r1 = env.make_handle(n, S, SEED)
    nb = L.pir_num_bits_for(n, 1, 0)
    total = args.warmup + args.steps
    timed = range(args.warmup, total)
    rng = np.random.default_rng(SEED)
    idxs = [int(x) for x in rng.integers(0, n, size=total)]
    qs = env.gen_keys(rng, nb, idxs)
    key = lambda i: qs[2 * i + (i & 1)]     # step i is answered for server (i & 1): both servers are timed and verified
    sp = env.sp
    d_out = torch.zeros((total + 1, S), dtype=torch.uint8, device="cuda")   # one answer per step (+ scratch row)
    comb = None
    EB = REGIONS * total          # combine slots: [r * total + i] step i of timed region r, then the e2e ring, a scratch slot, the pipelined ring
    SCRATCH = EB + RING
    if world > 1:
        # ring slots on rank 0's GPU: [0, REGIONS * total) one per device-timed step, [EB, EB + RING) re-used by the e2e loop,
        # one scratch slot; every rank's kernel XORs its partial straight into them (CUDA IPC mapping, NVLink)
        comb = env.sharding.open_combine_over_group(env.dist, env.local, EB + RING + 1 + PIPE_RING, S, owner=0)

    def step_device(i, slot=None):
        if world > 1:
            comb.accumulate(h, key(i), i if slot is None else slot, sp)
        else:
            capi.check(L.pir_answer_batch_device(h, C.byref(key(i)), 1, C.c_void_p(d_out[i if slot is None else slot].data_ptr()), sp))

    # ---- warm-up ----
    for i in range(args.warmup):
        step_device(i)
    env.barrier()

    # ---- device-resident timed region ----
    sampler = ClockSampler(env.local)
    if rank == 0:
        sampler.start()
    # The region of exactly `steps` steps (barrier + synchronize on both sides, CUDA events on the launching stream, max
    # over ranks) is run REGIONS times and the MEDIAN region is reported: one region lasts 3-18 ms, and a single host or
    # fabric hiccup inside it moved the N = 4 number by 10 % once in five runs.  All regions are listed in the line.
    region_ms = []
    for r in range(REGIONS):
        launches0 = L.pir_kernel_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        env.barrier()
        ev0.record(env.stream)
        for i in timed:
            step_device(i, slot=(r * total + i) if world > 1 else None)
        if world > 1 and rank == 0:
            comb.wait_device(r * total + total - 1, world, sp)      # the owner has seen all N arrivals of the last query
        ev1.record(env.stream)
        env.barrier()
        region_ms.append(ev0.elapsed_time(ev1))
        launches = L.pir_kernel_launch_count() - launches0 - (1 if world > 1 and rank == 0 else 0)
    region_ms = env.max_over_ranks(region_ms)
    dev_ms = sorted(region_ms)[REGIONS // 2]
    last = (REGIONS - 1) * total
    if world == 1:
        dev_ans = d_out[:total].cpu().numpy()
    elif rank == 0:    # (the answers of the last region; every region computes the same ones)
        dev_ans = np.stack([comb.wait_read(last + i, world, S) if i >= args.warmup else np.zeros(S, dtype=np.uint8) for i in range(total)])

    # ---- the same launches one at a time (CUDA events around every launch, no overlap between launches) ----
    capi.check(L.pir_db_set_kernel_timing(h, 1))
    for i in timed:
        step_device(i, slot=SCRATCH if world > 1 else total)
    kms, kcount = C.c_double(), C.c_uint64()
    capi.check(L.pir_db_get_kernel_timing(h, C.byref(kms), C.byref(kcount)))
    capi.check(L.pir_db_set_kernel_timing(h, 0))
    iso_ms = kms.value / max(1, kcount.value)
    # ---- the AES-only rate of this shard: BITS-mode launches of the same keys (no row is read) ----
    aes_ms = 0.0
    for i in timed:
        one = C.c_double()
        capi.check(L.pir_db_time_expansion(h, C.byref(key(i)), 1, C.byref(one)))
        aes_ms += one.value / len(timed)
    env.barrier()

    # ---- end-to-end timed region: host key in, host answer out, one query at a time ----
    e2e_ans = np.zeros((total, S), dtype=np.uint8)
    if world == 1:
        for i in range(min(3, args.warmup)):
            capi.check(L.pir_answer(h, C.byref(key(i)), e2e_ans[i].ctypes.data_as(C.c_void_p), S))
        env.barrier()
        t0 = time.perf_counter()
        for i in timed:
            capi.check(L.pir_answer(h, C.byref(key(i)), e2e_ans[i].ctypes.data_as(C.c_void_p), S))
        e2e_ms = (time.perf_counter() - t0) * 1e3
        # the same host call from CONC threads at once (a loaded server: one goroutine per query in flight).  A call holds
        # the handle's mutex only while it enqueues; the launches queue up on the GPU, overlap and look ahead.
        conc_ans = np.zeros((total, S), dtype=np.uint8)
        conc_err = []

        def conc_worker(t):
            try:
                for _ in range(CONC_PASSES):
                    for i in list(timed)[t::CONC]:
                        capi.check(L.pir_answer(h, C.byref(key(i)), conc_ans[i].ctypes.data_as(C.c_void_p), S))
            except Exception as exc:   # noqa: BLE001 (reported below)
                conc_err.append(exc)
        for phase in ("warm", "timed"):
            ths = [threading.Thread(target=conc_worker, args=(t,)) for t in range(CONC)]
            env.barrier()
            t0 = time.perf_counter()
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            conc_ms = (time.perf_counter() - t0) * 1e3
        if conc_err:
            raise SystemExit("bench.py: concurrent pir_answer failed: %s" % conc_err[0])
    else:
        box = Mailbox(env, C.sizeof(capi.PirQuery))

        def e2e_query(i, seq):
            if rank == 0:
                box.publish(key(i), seq)            # the key exists in rank 0's host memory only
                q = key(i)
            else:
                q = box.receive(seq)
            comb.accumulate(h, q, EB + (seq % RING), None)
            if rank == 0:                           # wait for the N arrivals, read, re-zero the slot: it is free again
                e2e_ans[i] = comb.wait_read_reset(EB + (seq % RING), world, S)
        seq = 0
        for i in range(min(3, args.warmup)):
            seq += 1
            e2e_query(i, seq)
        env.barrier()
        t0 = time.perf_counter()
        for i in timed:
            seq += 1
            e2e_query(i, seq)
        env.torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        env.barrier()
        # ---- the same end to end with PIPE_DEPTH queries in flight: rank 0 publishes query seq, every rank launches it,
        #      and rank 0 collects the answer of query seq - (PIPE_DEPTH - 1) (wait for its N arrivals, host copy, slot
        #      re-zeroed).  A combine slot / mailbox slot is re-used PIPE_RING queries later, after its answer was read.
        pipe_ans = np.zeros((total, S), dtype=np.uint8)
        pslot = lambda sq: EB + RING + 1 + (sq % PIPE_RING)

        def pipelined(idx_list, seq):
            pending = []
            for i in idx_list:
                seq += 1
                if rank == 0:
                    box.publish(key(i), seq)
                    q = key(i)
                else:
                    q = box.receive(seq)
                comb.accumulate(h, q, pslot(seq), None)
                if rank == 0:
                    pending.append((i, seq))
                    if len(pending) >= PIPE_DEPTH:
                        i0, s0 = pending.pop(0)
                        pipe_ans[i0] = comb.wait_read_reset(pslot(s0), world, S)
            for i0, s0 in pending:
                pipe_ans[i0] = comb.wait_read_reset(pslot(s0), world, S)
            env.torch.cuda.synchronize()
            return seq
        seq = pipelined(list(range(min(3, args.warmup))), seq)
        env.barrier()
        t0 = time.perf_counter()
        seq = pipelined(list(timed), seq)
        pipe_ms = (time.perf_counter() - t0) * 1e3
        env.barrier()
        box.close()
    clocks = sampler.stop() if rank == 0 else None

    # ---- N > 1: the single-process path (one process drives all N GPUs: what a Go server calls) ----
    sp_ms, sp_ans, spc_ms, spc_ans = None, None, None, None
    if world > 1:
        env.barrier()
        env.cpu_barrier()       # every rank's GPU is idle from here on; the others wait on the host (gloo), not in an NCCL kernel
        if rank == 0:
            sp_ans = np.zeros((total, S), dtype=np.uint8)
            with env.sharding.Server(list(range(world)), n, S) as srv:
                srv.fill_synthetic(SEED)
                for i in range(min(3, args.warmup)):
                    sp_ans[i] = srv.answer(key(i))
                t0 = time.perf_counter()
                for i in timed:
                    sp_ans[i] = srv.answer(key(i))
                sp_ms = (time.perf_counter() - t0) * 1e3
                # the same call from CONC host threads at once (a Go server with one goroutine per query in flight)
                spc_ans = np.zeros((total, S), dtype=np.uint8)
                spc_err = []

                def spc_worker(t):
                    try:
                        for _ in range(CONC_PASSES):
                            for i in list(timed)[t::CONC]:
                                spc_ans[i] = srv.answer(key(i))
                    except Exception as exc:   # noqa: BLE001 (reported below)
                        spc_err.append(exc)
                for phase in ("warm", "timed"):
                    ths = [threading.Thread(target=spc_worker, args=(t,)) for t in range(CONC)]
                    t0 = time.perf_counter()
                    for th in ths:
                        th.start()
                    for th in ths:
                        th.join()
                    spc_ms = (time.perf_counter() - t0) * 1e3
                if spc_err:
                    raise SystemExit("bench.py: concurrent pir_server_answer failed: %s" % spc_err[0])
        env.cpu_barrier()
        env.barrier()

    if world > 1:
        conc_ans, conc_ms = None, None
        pipe_ms = env.max_over_ranks([pipe_ms])[0]
    else:
        pipe_ans, pipe_ms = None, None
    e2e_ms, iso_ms, aes_ms = env.max_over_ranks([e2e_ms, iso_ms, aes_ms])

    # ---- parity: every timed answer against the oracle's bytes (outside the timed regions) ----
    checked, mean_sel = 0, None
    if rank == 0:
        want, mean_sel = oracle_answers(env, n, S, SEED, [key(i) for i in timed])
        for k, i in enumerate(timed):
            for name, got in (("device-timed", dev_ans), ("e2e", e2e_ans), ("single-process", sp_ans), ("concurrent e2e", conc_ans),
                              ("pipelined e2e", pipe_ans), ("concurrent single-process", spc_ans)):
                if got is not None and not np.array_equal(got[i], want[k]):
                    raise SystemExit("bench.py: %s answer of step %d (server %d) differs from the oracle's bytes -- "
                                     "refusing to report a wrong path" % (name, i, i & 1))
            checked += 1
        # and the two servers' answers recover the slot (query.go:278-297) for the first timed key
        i = args.warmup
        other = oracle_answers(env, n, S, SEED, [qs[2 * i + ((i & 1) ^ 1)]])[0][0]
        import oracle
        if not np.array_equal(other ^ dev_ans[i], oracle.synth_fill(SEED, idxs[i], 1, S)[0]):
            raise SystemExit("bench.py: recovered slot differs from the database slot")

    line = None
    if rank == 0:
        db_bytes = n * S
        ms_step = dev_ms / args.steps
        e2e_step = e2e_ms / args.steps
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        shard_bytes = (r1 - r0) * S
        achieved = shard_bytes / (ms_step * 1e-3) / 1e9
        phys = mean_sel / world * S          # selected rows of one shard, on average
        even = sum(1 for i in timed if not key(i).final_cw & 1)
        line = {
            "metric": METRIC, "value": db_bytes / (ms_step * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(), "sharding": "rows/%d" % world,
                       "l2": "per-GPU shard (%d MiB) exceeds the 126 MB L2; every step scans all of it" % (shard_bytes >> 20),
                       "combine": "none" if world == 1 else "every shard's kernel XORs its %d-byte partial into rank 0's ring slot "
                                  "(red.global.xor over a CUDA IPC mapping / NVLink); no NCCL in the data path" % S,
                       "queries": "distinct seeded key per step, server share alternating; FinalCW even / odd in the timed steps: %d / %d "
                                  "(a coin flip per key; the kernel has a code path per parity)" % (even, args.steps - even),
                       "value_regions_ms": [round(x, 4) for x in region_ms],
                       "value_covers": "median of 3 timed regions of exactly `steps` steps each (all listed in value_regions_ms; each with its own "
                                       "barriers, CUDA events and max over ranks); a region = back-to-back single-query launches on one stream per GPU, consecutive launches overlapping "
                                       "(programmatic dependent launch)" + ("" if world == 1 else
                                       "; rank 0's event pair ends after the device-side wait for the last query's %d arrivals; max over ranks" % world)},
            "queries_per_s": 1e3 / ms_step,
            "e2e_one_at_a_time": {"value": db_bytes / (e2e_step * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_step,
                    "h2d_bytes_per_step": C.sizeof(capi.PirQuery) * world, "d2h_bytes_per_step": S,
                    "queries_per_s": 1e3 / e2e_step,
                    "api": "pir_answer (C ABI, host key in / host answer out, one launch + one stream sync per query)" if world == 1 else
                           "per query, inside the timed loop: key fanned out from rank 0's host memory to the other processes "
                           "(shared-memory mailbox), pir_answer_accumulate_device on every rank, pir_combine_wait_read_reset on rank 0 "
                           "(device-side wait for %d arrivals, combined answer to the host, slot re-zeroed; ring of %d slots)" % (world, RING)},
            "gpu_launches": int(launches),
            "parity_checked": checked,
            "parity": "every timed answer of both timed regions%s == oracle bytes (bits: oracle.expand_tree, rows: oracle.answers_over_synth); "
                      "servers 0 and 1 alternate; recover-to-slot checked" % ("" if world == 1 else " and of the single-process path"),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": phys, "traffic_source": "selected rows only (unselected rows are never fetched, like db.go:93): mean popcount of the "
                                                             "timed keys' selection bits x row bytes, per shard; ncu dram__bytes of one launch: profiles/",
                         "kernel": "answer_kernel", "kernel_ms": ms_step, "kernel_ms_isolated": iso_ms,
                         "kernel_ms_note": "kernel_ms = timed region / launches (the region is exactly `steps` launches of this kernel per GPU, "
                                           "overlapping at their edges); isolated = CUDA events around every launch, launches serialised",
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650",
                         "algorithmic_bytes_per_launch": shard_bytes,
                         "lookup": {"achieved": 3.0 * (r1 - r0) / (ms_step * 1e-3), "peak": 3.0 * (r1 - r0) / (aes_ms * 1e-3),
                                    "unit": "AES blocks/s (3 per row)", "frac": aes_ms / ms_step, "aes_only_ms": aes_ms,
                                    "peak_source": "this run: BITS-mode launches of the same keys on the same shard (same tree expansion and "
                                                   "work split, no database row read), CUDA events, serialised; the T-table AES is bound by the "
                                                   "shared-memory lookup pipe (ncu: l1tex data pipe > 90 %), so this is the bound that binds"},
                         "dram": {"achieved": phys / (ms_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                  "frac": phys / (ms_step * 1e-3) / 1e9 / peak}},
            "clocks": clocks,
        }
        if conc_ms is not None:
            per = conc_ms / (args.steps * CONC_PASSES)
            line["e2e"] = {"value": db_bytes / (per * 1e-3) / 1e9, "unit": UNIT, "ms_per_query": per, "queries_per_s": 1e3 / per,
                                      "ms_per_step": per, "h2d_bytes_per_step": C.sizeof(capi.PirQuery), "d2h_bytes_per_step": S,
                                      "in_flight": CONC, "threads": CONC, "queries": args.steps * CONC_PASSES,
                                      "note": "throughput with queries in flight, like `value` (back-to-back launches); every query's key comes "
                                              "from host memory and its answer goes to host memory inside the timed region; one query at a time: "
                                              "e2e_one_at_a_time",
                                      "api": "pir_answer (host key in / host answer out) from %d host threads at once on ONE handle: each call "
                                             "enqueues under the handle's mutex and waits for its own event with the mutex released" % CONC}
        if pipe_ms is not None:
            per = pipe_ms / args.steps
            line["e2e"] = {"value": db_bytes / (per * 1e-3) / 1e9, "unit": UNIT, "ms_per_query": per, "queries_per_s": 1e3 / per,
                                     "ms_per_step": per, "h2d_bytes_per_step": C.sizeof(capi.PirQuery) * world, "d2h_bytes_per_step": S,
                                     "in_flight": PIPE_DEPTH,
                                     "note": "throughput with queries in flight, like `value` (back-to-back launches); every query's key comes "
                                             "from rank 0's host memory and its combined answer goes to rank 0's host memory inside the timed "
                                             "region; one query at a time: e2e_one_at_a_time",
                                     "api": "as e2e_one_at_a_time (key fanned out from rank 0's host memory, pir_answer_accumulate_device on every rank, "
                                            "pir_combine_wait_read_reset on rank 0), with %d queries in flight: rank 0 collects answer k - %d "
                                            "after launching query k; combine ring of %d slots re-used inside the loop" % (
                                                PIPE_DEPTH, PIPE_DEPTH - 1, PIPE_RING)}
        if sp_ms is not None:
            line["e2e_single_process"] = {"value": db_bytes / (sp_ms / args.steps * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": sp_ms / args.steps,
                                          "api": "pir_server_answer: ONE process (rank 0) drives all %d GPUs, host key in / host answer out; the other "
                                                 "ranks idle meanwhile" % world}
            per = spc_ms / (args.steps * CONC_PASSES)
            line["e2e_single_process_concurrent"] = {"value": db_bytes / (per * 1e-3) / 1e9, "unit": UNIT, "ms_per_query": per,
                                                     "queries_per_s": 1e3 / per, "threads": CONC,
                                                     "api": "pir_server_answer from %d host threads at once on ONE server handle" % CONC}
    # ---- batches: >= 32 queries in flight (queries/s) ----
    NQ = 32
    qb = (capi.PirQuery * NQ)(*[qs[2 * (k % total) + (k & 1)] for k in range(NQ)])
    if world == 1:
        d_b = torch.zeros((NQ, S), dtype=torch.uint8, device="cuda")
        capi.check(L.pir_answer_batch_device(h, qb, NQ, C.c_void_p(d_b.data_ptr()), sp))
        env.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(env.stream)
        for _ in range(2):
            capi.check(L.pir_answer_batch_device(h, qb, NQ, C.c_void_p(d_b.data_ptr()), sp))
        e1.record(env.stream)
        env.barrier()
        if rank == 0:
            got = d_b.cpu().numpy()
            want, _ = oracle_answers(env, n, S, SEED, [qb[k] for k in range(8)])
            if not np.array_equal(got[:8], want):
                raise SystemExit("bench.py: batch answers differ from the oracle's bytes")
            line["queries_per_s_batch"] = {"value": 2e3 * NQ / e0.elapsed_time(e1), "queries_in_flight": NQ, "parity_checked": 8,
                                           "api": "pir_answer_batch_device: one fused launch per query alternating between two streams"}
    if comb is not None:
        env.barrier()
        if rank != 0:
            comb.close()   # peers unmap before the owner frees
        env.dist.barrier()
        if rank == 0:
            comb.close()
    L.pir_db_destroy(h)
    return line


# ------------------------------------------------------------------------------------------------
# the other BASELINE configs, on the shipped kernels, each verified against the oracle
# ------------------------------------------------------------------------------------------------
def other_config(env, name, log_rows, S, queries=1, keyword=False, steps=6, warm=3, seed=77):
    """One JSON object for BASELINE configs[1] / [3] / [4]: device-timed ms per step, GB/s, queries/s, both rooflines,
    and how many answers were compared with the oracle's bytes."""
    np, torch, capi, L = env.np, env.torch, env.capi, env.L
    world, rank, sp = env.world, env.rank, env.sp
    n = 1 << log_rows
    h, r0, r1 = env.make_handle(n, S, seed)
    kws = None
    if keyword:
        kws = np.ascontiguousarray(keywords_for(r0, r1, log_rows))
        capi.check(L.pir_db_set_keywords(h, kws.ctypes.data_as(C.POINTER(C.c_uint64)), kws.size))
    nb = L.pir_num_bits_for(n, 1, 1 if keyword else 0)
    total = warm + steps
    rng = np.random.default_rng(seed)
    idxs = [int(x) for x in rng.integers(0, n, size=total * queries)]
    targets = [int(keywords_for(i, i + 1, log_rows)[0]) for i in idxs] if keyword else idxs
    qs = env.gen_keys(rng, nb, targets, keyword)
    key = lambda k: qs[2 * k + (k & 1)]
    comb = None
    if world > 1:
        comb = env.sharding.open_combine_over_group(env.dist, env.local, total * queries + 1, (S + 15) // 16 * 16, owner=0)
    d_out = torch.zeros((total * queries, S), dtype=torch.uint8, device="cuda")

    def step(i):
        if world > 1:
            for k in range(i * queries, (i + 1) * queries):
                comb.accumulate(h, key(k), k, sp)
        elif queries == 1:
            capi.check(L.pir_answer_batch_device(h, C.byref(key(i)), 1, C.c_void_p(d_out[i].data_ptr()), sp))
        else:
            arr = (capi.PirQuery * queries)(*[key(k) for k in range(i * queries, (i + 1) * queries)])
            capi.check(L.pir_answer_batch_device(h, arr, queries, C.c_void_p(d_out[i * queries].data_ptr()), sp))
    for i in range(warm):
        step(i)
    env.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(env.stream)
    for i in range(warm, total):
        step(i)
    if world > 1 and rank == 0:
        comb.wait_device(total * queries - 1, world, sp)
    ev1.record(env.stream)
    env.barrier()
    ms = ev0.elapsed_time(ev1) / steps
    aes_ms = None
    if not keyword:
        one = C.c_double()
        capi.check(L.pir_db_time_expansion(h, C.byref(key(warm * queries)), 2, C.byref(one)))
        aes_ms = one.value
    ms, aes = env.max_over_ranks([ms, aes_ms or 0.0])
    first = warm * queries
    if world == 1:
        ans = d_out.cpu().numpy()
    elif rank == 0:
        ans = np.zeros((total * queries, S), dtype=np.uint8)
        for k in range(first, total * queries):
            ans[k] = comb.wait_read(k, world, S)
    # ---- parity ----
    import oracle
    checked, how = 0, ""
    if not keyword:
        if rank == 0:
            ks = list(range(first, total * queries))
            want, _ = oracle_answers(env, n, S, seed, [key(k) for k in ks])
            for j, k in enumerate(ks):
                if not np.array_equal(ans[k], want[j]):
                    raise SystemExit("bench.py: %s: answer %d differs from the oracle's bytes" % (name, k))
            checked, how = len(ks), "every timed answer == oracle bytes (expand_tree + answers_over_synth)"
    else:
        # Keyword queries: the oracle walks 32 levels per row, so the bits are compared on a 2^18-row window of every
        # shard, and the answer is re-derived by the oracle's row XOR over the GPU's bits of the whole shard (partials
        # XORed over the ranks); for the other timed keys the two servers' answers must recover the slot.
        k = first
        q = key(k)
        bits = np.zeros(r1 - r0, dtype=np.uint8)
        capi.check(L.pir_expand_bits(h, C.byref(q), bits.ctypes.data_as(C.POINTER(C.c_uint8)), r1 - r0))
        w = min(1 << 18, r1 - r0)
        ob = oracle.expand_shared_query(oracle.query_from_bytes(q.to_bytes()), n, keywords=kws[:w], row_begin=0, n_rows=w, nthreads=env.cores)
        ok = bool(np.array_equal(bits[:w], ob))
        part = oracle.answers_over_synth(seed, S, 1, [bits], row_begin=r0, nthreads=env.cores)[0]
        if world > 1:
            parts = [None] * world
            env.dist.all_gather_object(parts, (ok, part.tobytes()))
        else:
            parts = [(ok, part.tobytes())]
        if rank == 0:
            if not all(p[0] for p in parts):
                raise SystemExit("bench.py: %s: keyword selection bits differ from the oracle's" % name)
            want = np.zeros(S, dtype=np.uint8)
            for p in parts:
                want ^= np.frombuffer(p[1], dtype=np.uint8)
            if not np.array_equal(ans[k], want):
                raise SystemExit("bench.py: %s: keyword answer differs from the oracle's row XOR" % name)
            checked, how = 1, ("1 timed answer: bits == oracle per-row Evaluate2P on a 2^18-row window of every shard, answer == oracle row XOR "
                               "over the bits; all timed answers: the two servers recover the slot")
        # recover-to-slot for every timed key: answer the other server's share now
        for k in range(first, total * queries):
            other = qs[2 * k + ((k & 1) ^ 1)]
            if world > 1:
                comb.accumulate(h, other, total * queries, sp)
                env.barrier()
                if rank == 0:
                    o = comb.wait_read_reset(total * queries, world, S)
                env.barrier()
            else:
                capi.check(L.pir_answer_batch_device(h, C.byref(other), 1, C.c_void_p(d_out[0].data_ptr()), sp))
                env.torch.cuda.synchronize()
                o = d_out[0].cpu().numpy()
            if rank == 0 and not np.array_equal(o ^ ans[k], oracle.synth_fill(seed, idxs[k], 1, S)[0]):
                raise SystemExit("bench.py: %s: keyword answers do not recover slot %d" % (name, idxs[k]))
    out = None
    if rank == 0:
        peak = 6650.0
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        shard_rows = (r1 - r0)
        gbs = n * S * queries / (ms * 1e-3) / 1e9
        out = {"config": name, "workload": workload_name(log_rows, S, queries, keyword), "n_gpus": world, "steps": steps, "warmup": warm,
               "ms_per_step": ms, "value": gbs, "unit": UNIT + (" (effective: one pass per query)" if queries > 1 else ""),
               "queries_per_s": 1e3 * queries / ms,
               "roofline": {"hbm": {"achieved": shard_rows * S * queries / (ms * 1e-3) / 1e9, "peak": peak,
                                    "frac": shard_rows * S * queries / (ms * 1e-3) / 1e9 / peak}},
               "parity_checked": checked, "parity": how}
        if queries > 1:
            out["roofline"]["hbm"]["note"] = "algorithmic bytes: every query scans the database; the two-phase batch streams it once per batch"
            out["roofline"]["hbm_per_pass"] = {"achieved": shard_rows * S / (ms * 1e-3) / 1e9, "frac": shard_rows * S / (ms * 1e-3) / 1e9 / peak}
        if aes_ms is not None:
            out["roofline"]["lookup"] = {"achieved": 3.0 * shard_rows * queries / (ms * 1e-3), "peak": 3.0 * shard_rows / (aes * 1e-3),
                                         "unit": "AES blocks/s (3 per row)", "frac": aes * queries / ms, "aes_only_ms": aes}
        else:
            out["roofline"]["walk"] = {"rows_per_s": shard_rows / (ms * 1e-3),
                                       "note": "32-level DPF at Keywords[row] with prefix sharing between consecutive rows; no AES-only reference launch"}
    if comb is not None:
        env.barrier()
        if rank != 0:
            comb.close()
        env.dist.barrier()
        if rank == 0:
            comb.close()
    L.pir_db_destroy(h)
    del d_out
    torch.cuda.empty_cache()
    return out


def gpu_arm(args):
    env = Env(args)
    line = headline(env, args)
    if OTHER:
        others = []
        if env.world == 1:
            others.append(other_config(env, "configs[1]", 20, 256, steps=10, warm=3, seed=31))
            others.append(other_config(env, "configs[3]", 22, 1024, queries=64, steps=3, warm=2, seed=33))
        others.append(other_config(env, "configs[4] index", 28, 32, steps=4, warm=2, seed=35))
        others.append(other_config(env, "configs[4] keyword", 28, 32, keyword=True, steps=3, warm=2, seed=37))
        if env.rank == 0:
            line["other_configs"] = others
    if env.rank == 0:
        if env.world == 1:
            gbs, ms, cores, sample = cpu_reference_run(3, 1)
            line["cpu_baseline"] = {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if env.world > 1:
        env.dist.destroy_process_group()


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
