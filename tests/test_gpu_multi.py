"""GPU tests of the multi-GPU combine: pir_server (one process, all shards) and
pir_combine (one process per GPU, CUDA IPC).  On a 1-GPU box the shards share device 0, which exercises
the same sharding arithmetic, accumulate-without-zeroing kernels and arrival counters; with >= 2 GPUs the
peer-store path over NVLink is covered too."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import oracle
from pir_b200 import _capi, sharding
from tests.helpers import Handle, make_queries

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices(gpu_lib, n):
    have = gpu_lib.pir_device_count()
    return [i % have for i in range(n)]


@pytest.mark.parametrize("n_shards,db_size,slot_bytes,W", [
    (1, 5000, 32, 1), (2, 1 << 14, 256, 1), (3, 12345, 64, 2), (4, 3000, 100, 3), (8, 1 << 15, 32, 1),
    (2, 1 << 12, 1024, 4), (5, 7, 16, 1),
])
def test_server_answer_matches_oracle(gpu_lib, n_shards, db_size, slot_bytes, W):
    data = oracle.synth_fill(5, 0, db_size, slot_bytes)
    rows = db_size // W
    idx = (rows * 2) // 3
    (o0, o1), (q0, q1) = make_queries(11, db_size, W, idx)
    with sharding.Server(_devices(gpu_lib, n_shards), db_size, slot_bytes, shard_align=W) as srv:
        assert srv.num_shards == n_shards
        srv.fill_synthetic(5)
        a0, a1 = srv.answer(q0), srv.answer(q1)
        assert np.array_equal(a0, oracle.answer(o0, data, db_size, slot_bytes))
        assert np.array_equal(a1, oracle.answer(o1, data, db_size, slot_bytes))
        assert np.array_equal(a0 ^ a1, data[idx * W:(idx + 1) * W].reshape(-1))
        both = srv.answer_batch([q0, q1, q0])
        assert np.array_equal(both[0], a0) and np.array_equal(both[1], a1) and np.array_equal(both[2], a0)


@pytest.mark.parametrize("n_shards,db_size,slot_bytes", [(1, 900, 32), (3, 5000, 32), (8, 1 << 14, 40)])
def test_server_keyword_query_over_shards(gpu_lib, n_shards, db_size, slot_bytes):
    """BASELINE configs[4]'s keyword half at test size: a keyword query (db.go:119-121,132-135) over a
    row-sharded database; every shard evaluates the 32-level DPF at its own rows' keywords."""
    rng = np.random.default_rng(21)
    data = rng.integers(0, 256, size=(db_size, slot_bytes), dtype=np.uint8)
    kws = np.sort(rng.choice(1 << 32, size=db_size, replace=False).astype(np.uint64))
    target = (db_size * 5) // 7
    (o0, o1), (q0, q1) = make_queries(13, db_size, 1, target, is_keyword=True, keyword=int(kws[target]))
    with sharding.Server(_devices(gpu_lib, n_shards), db_size, slot_bytes) as srv:
        srv.upload(data)
        with pytest.raises(_capi.PirError):
            srv.answer(q0)                      # no keywords yet
        with pytest.raises(_capi.PirError):
            srv.set_keywords(kws[:-1])          # one keyword per slot
        srv.set_keywords(kws)
        a0, a1 = srv.answer(q0), srv.answer(q1)
        assert np.array_equal(a0, oracle.answer(o0, data, db_size, slot_bytes, keywords=kws))
        assert np.array_equal(a1, oracle.answer(o1, data, db_size, slot_bytes, keywords=kws))
        assert np.array_equal(a0 ^ a1, data[target])


@pytest.mark.parametrize("n_shards,db_size,slot_bytes,W", [(2, 4096, 32, 2), (3, 5001, 48, 3), (8, 1 << 14, 32, 4), (4, 900, 272, 2)])
def test_server_keyword_query_with_a_group_size_over_shards(gpu_lib, n_shards, db_size, slot_bytes, W):
    """Keyword query with GroupSize > 1 on a row-sharded database: row r evaluates Keywords[r] (db.go:132-135), which for every
    shard but the first is a keyword of an EARLIER shard's slots -- pir_server_set_keywords hands each shard its rows' keywords."""
    rng = np.random.default_rng(23)
    data = rng.integers(0, 256, size=(db_size, slot_bytes), dtype=np.uint8)
    kws = np.sort(rng.choice(1 << 32, size=db_size, replace=False).astype(np.uint64))
    rows = db_size // W
    for target in (rows - 1, (rows * 5) // 7, 0):
        (o0, o1), (q0, q1) = make_queries(14, db_size, W, target, is_keyword=True, keyword=int(kws[target]))
        with sharding.Server(_devices(gpu_lib, n_shards), db_size, slot_bytes, shard_align=W) as srv:
            srv.upload(data)
            srv.set_keywords(kws)
            a0, a1 = srv.answer(q0), srv.answer(q1)
            assert np.array_equal(a0, oracle.answer(o0, data, db_size, slot_bytes, keywords=kws))
            assert np.array_equal(a1, oracle.answer(o1, data, db_size, slot_bytes, keywords=kws))
            assert np.array_equal(a0 ^ a1, data[target * W:(target + 1) * W].reshape(-1))
            both = srv.answer_batch([q1, q0])
            assert np.array_equal(both[0], a1) and np.array_equal(both[1], a0)
            if target == 0 and n_shards > 1:
                # group size 1 still works on the same server; another group size that divides the shards does not have its
                # rows' keywords on the later shards and says so
                (p0, _), (g0, _) = make_queries(15, db_size, 1, db_size - 1, is_keyword=True, keyword=int(kws[db_size - 1]))
                assert np.array_equal(srv.answer(g0), oracle.answer(p0, data, db_size, slot_bytes, keywords=kws))
                if W == 4:
                    (_, _), (h0, _) = make_queries(16, db_size, 2, 5, is_keyword=True, keyword=int(kws[5]))
                    with pytest.raises(_capi.PirError) as e:
                        srv.answer(h0)
                    assert "pir_db_set_row_keywords" in str(e.value)


def test_shard_handle_row_keywords(gpu_lib):
    """pir_db_set_row_keywords on a bare shard handle: ExpandSharedQuery of a keyword query with GroupSize 3 on the middle
    shard equals the oracle's bits for those rows; without the row keywords the call fails with a clear error."""
    rng = np.random.default_rng(29)
    db_size, S, W = 3000, 64, 3
    kws = np.sort(rng.choice(1 << 32, size=db_size, replace=False).astype(np.uint64))
    (o0, _), (q0, _) = make_queries(17, db_size, W, 444, is_keyword=True, keyword=int(kws[444]))
    with Handle(db_size, S, slot_begin=999, slot_count=1002) as h:
        h.fill(3)
        h.set_keywords(kws[999:999 + 1002])
        r0, n = h.rows(W)
        assert (r0, n) == (333, 334)
        with pytest.raises(_capi.PirError) as e:
            h.expand_bits(q0)
        assert "pir_db_set_row_keywords" in str(e.value)
        with pytest.raises(_capi.PirError):
            h.set_row_keywords(W, kws[r0:r0 + n - 1])           # one keyword per row of the shard
        h.set_row_keywords(W, kws[r0:r0 + n])
        want = oracle.expand_shared_query(o0, db_size, keywords=kws, row_begin=r0, n_rows=n)
        assert np.array_equal(h.expand_bits(q0), want) and want[444 - r0] != oracle.expand_shared_query(
            make_queries(17, db_size, W, 444, is_keyword=True, keyword=int(kws[444]))[0][1], db_size, keywords=kws, row_begin=r0, n_rows=n)[444 - r0]
        data = oracle.synth_fill(3, 0, db_size, S)
        part = np.zeros(W * S, dtype=np.uint8)
        for r in np.nonzero(want)[0]:
            part ^= data[(r0 + r) * W:(r0 + r + 1) * W].reshape(-1)
        assert np.array_equal(h.answer(q0), part)
        # a different group size replaces the copy
        h.set_row_keywords(1, kws[999:999 + 1002])
        with pytest.raises(_capi.PirError):
            h.expand_bits(q0)


def test_server_upload_routes_to_shards(gpu_lib):
    rng = np.random.default_rng(3)
    db_size, S = 1000, 48
    data = rng.integers(0, 256, size=(db_size, S), dtype=np.uint8)
    (o0, _), (q0, _) = make_queries(4, db_size, 1, 777)
    with sharding.Server(_devices(gpu_lib, 3), db_size, S) as srv:
        srv.upload(data[:400], 0)
        srv.upload(data[400:], 400)   # straddles shard boundaries
        assert np.array_equal(srv.answer(q0), oracle.answer(o0, data, db_size, S))


@pytest.mark.parametrize("direct", [0, 1])
def test_combine_slots_single_process(gpu_lib, direct):
    """Two shard handles accumulate into ring slots of one combine buffer; the owner waits for both.  direct = 0: the CTAs
    fold into the handle's local accumulator and the launch's last CTA forwards the sum to the slot (shipped);
    direct = 1: every CTA XORs into the slot itself (round 1)."""
    n, S = 1 << 13, 256
    data = oracle.synth_fill(9, 0, n, S)
    comb = sharding.Combine(0, 4, S)
    try:
        with Handle(n, S, 0, n // 2) as h0, Handle(n, S, n // 2, n // 2) as h1:
            h0.set_option("direct_combine", direct)
            h1.set_option("direct_combine", direct)
            h0.fill(9)
            h1.fill(9)
            for rep in range(2):  # second round re-uses the slots after a reset
                for slot, idx in enumerate((5, 4097, 8000, 4095)):
                    (o0, _), (q0, _) = make_queries(20 + slot + 10 * rep, n, 1, idx)
                    comb.accumulate(h0.h, q0, slot)
                    comb.accumulate(h1.h, q0, slot)
                    got = comb.wait_read(slot, 2, S)
                    assert np.array_equal(got, oracle.answer(o0, data, n, S)), (rep, slot)
                comb.reset()
    finally:
        comb.close()


def test_combine_wait_read_reset_frees_the_slot(gpu_lib):
    """pir_combine_wait_read_reset: one launch waits for the arrivals, moves the answer to the host and zeroes slot and
    counter, so the same slot serves the next query at once (the end-to-end loop of bench.py: ring of 4 slots)."""
    n, S = 1 << 13, 256
    data = oracle.synth_fill(9, 0, n, S)
    comb = sharding.Combine(0, 2, S)
    try:
        with Handle(n, S, 0, n // 2) as h0, Handle(n, S, n // 2, n // 2) as h1:
            h0.fill(9)
            h1.fill(9)
            for k in range(12):
                (o0, o1), (q0, q1) = make_queries(300 + k, n, 1, (k * 811) % n)
                o, q = ((o0, q0), (o1, q1))[k & 1]
                slot = k & 1
                comb.accumulate(h0.h, q, slot)
                comb.accumulate(h1.h, q, slot)
                got = comb.wait_read_reset(slot, 2, S)
                assert np.array_equal(got, oracle.answer(o, data, n, S)), k
            with pytest.raises(_capi.PirError):
                comb.wait_read_reset(0, 1, S, timeout_ms=20)      # nothing pending: times out, the slot stays clean
            assert not comb.wait_read(0, 0, S).any() and not comb.wait_read(1, 0, S).any()
    finally:
        comb.close()


def test_combine_wait_times_out(gpu_lib):
    comb = sharding.Combine(0, 1, 64)
    try:
        with pytest.raises(_capi.PirError):
            comb.wait_read(0, 1, 64, timeout_ms=20)
        # the flag is cleared: a satisfied wait works afterwards
        assert np.array_equal(comb.wait_read(0, 0, 64, timeout_ms=20), np.zeros(64, dtype=np.uint8))
    finally:
        comb.close()


def _ipc_worker(rank, world, port, n, S, idx, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pir_b200 import sharding as sh
        from tests.helpers import Handle as H, make_queries as mq
        import oracle as orc
        comb = sh.open_combine_over_group(dist, rank, 2, S)
        begin, count = sh.shard_slots(n, 1, rank, world)
        with H(n, S, begin, count, device=rank) as h:
            h.fill(13)
            (o0, o1), (q0, q1) = mq(77, n, 1, idx)
            comb.accumulate(h.h, q0, 0)
            comb.accumulate(h.h, q1, 1)
            if rank == 0:
                a0 = comb.wait_read(0, world, S)
                a1 = comb.wait_read(1, world, S)
                data = orc.synth_fill(13, 0, n, S)
                ok = bool(np.array_equal(a0, orc.answer(o0, data, n, S)) and
                          np.array_equal(a1, orc.answer(o1, data, n, S)) and
                          np.array_equal(a0 ^ a1, data[idx]))
                q.put(ok)
            dist.barrier()
        if rank != 0:
            comb.close()   # peers unmap before the owner frees
        dist.barrier()
        if rank == 0:
            comb.close()
    finally:
        dist.destroy_process_group()


def test_combine_over_cuda_ipc_two_gpus(gpu_lib):
    """One process per GPU: rank 1's kernel XORs into rank 0's memory through a CUDA IPC mapping."""
    if gpu_lib.pir_device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ipc_worker, args=(r, 2, port, 1 << 14, 256, 9999, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("rank hung")
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_server_batch_two_phase_over_shards(gpu_lib):
    """pir_server_answer_batch with wide rows: every shard runs the bit-vector + single-pass scan path and
    XORs into the shared accumulator."""
    db_size, S, nq = 1 << 13, 1024, 6
    data = oracle.synth_fill(8, 0, db_size, S)
    pairs = [make_queries(50 + i, db_size, 1, (i * 1237) % db_size) for i in range(nq)]
    qs = [p[1][0] for p in pairs]
    with sharding.Server(_devices(gpu_lib, 3), db_size, S) as srv:
        srv.set_option("batch_mode", 2)
        srv.fill_synthetic(8)
        got = srv.answer_batch(qs)
    for i, p in enumerate(pairs):
        assert np.array_equal(got[i], oracle.answer(p[0][0], data, db_size, S)), i


@pytest.mark.parametrize("n_shards,db_size,slot_bytes,W", [(3, 1 << 16, 256, 1), (8, 50000, 48, 2), (2, 1 << 15, 1024, 1)])
def test_server_answer_from_many_threads(gpu_lib, n_shards, db_size, slot_bytes, W):
    """pir_server_answer from 10 threads at once (more than the server's ring of 8 in-flight slots): one launch per
    shard into a ring slot, one wait+read+reset launch on the owner, the server's mutex released while a caller waits.
    Every answer must equal the oracle's, also when batch calls (the other path) run in between."""
    import threading
    data = oracle.synth_fill(61, 0, db_size, slot_bytes)
    rows = db_size // W
    rng = np.random.default_rng(n_shards)
    pairs = [make_queries(6000 + k, db_size, W, int(rng.integers(0, rows))) for k in range(15)]
    keys = [(p[0][s], p[1][s]) for p in pairs for s in (0, 1)]
    want = [oracle.answer(o, data, db_size, slot_bytes, use_tree=True) for o, _ in keys]
    errors = []
    with sharding.Server(_devices(gpu_lib, n_shards), db_size, slot_bytes, shard_align=W) as srv:
        srv.set_option("direct_combine", n_shards == 8)   # (both combine variants are covered)
        srv.fill_synthetic(61)

        def worker(t):
            try:
                for rep in range(3):
                    for k in range(t, len(keys), 10):
                        if not np.array_equal(srv.answer(keys[k][1]), want[k]):
                            errors.append((t, rep, k))
                    if t == 0:
                        got = srv.answer_batch([keys[1][1], keys[2][1]])
                        if not (np.array_equal(got[0], want[1]) and np.array_equal(got[1], want[2])):
                            errors.append(("batch", rep))
            except Exception as exc:   # noqa: BLE001
                errors.append((t, repr(exc)))
        ths = [threading.Thread(target=worker, args=(t,)) for t in range(10)]
        for th in ths:
            th.start()
        for th in ths:
            th.join(timeout=180)
        assert not any(th.is_alive() for th in ths), "a thread hung"
        # a wrong key is rejected and the ring keeps working
        bad = _capi.PirQuery.from_bytes(keys[0][1].to_bytes())
        bad.t_init = 9
        with pytest.raises(_capi.PirError):
            srv.answer(bad)
        assert np.array_equal(srv.answer(keys[3][1]), want[3])
    assert not errors, errors[:5]
