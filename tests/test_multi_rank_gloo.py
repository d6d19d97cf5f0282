"""world_size-2 (and 3) `gloo` test of the multi-GPU host logic on CPU: row sharding + XOR combine.

Each rank plays one GPU of a row-sharded server.  No CUDA device exists here, so the per-shard partial
answer comes from the oracle restricted to the rank's rows (the checker standing in for the kernel —
the GPU version of this test is tests/test_gpu_parity.py::test_sharded_handles_combine); what is under
test is the product's sharding arithmetic (pir_b200/sharding.py), the gather over a real process group
and pir_xor_combine."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, S, W, idx, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import oracle
    from pir_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o0, o1 = oracle.gen_two_server(oracle.tape(31), oracle.num_bits_for(n, W), idx, 1)
        begin, count = sharding.shard_slots(n, W, rank, world)
        r0, nr = sharding.shard_rows(n, W, rank, world)
        assert begin == r0 * W
        shard = oracle.synth_fill(77, begin, count, S)       # this rank's slots only
        combined = []
        for o in (o0, o1):
            o.group_size = W
            bits = oracle.expand_tree(o, n, r0, nr)          # this rank's subtree of the DPF domain
            part = oracle.answer_with_bits(shard, S, W, bits)
            combined.append(sharding.gather_and_combine(part, dist, dst=0))
        if rank == 0:
            full = oracle.synth_fill(77, 0, n, S)
            ok = True
            for o, c in zip((o0, o1), combined):
                ok &= bool(np.array_equal(c, oracle.answer(o, full, n, S)))
            ok &= bool(np.array_equal(combined[0] ^ combined[1], full[idx * W:(idx + 1) * W].reshape(-1)))
            q.put(ok)
        else:
            assert combined == [None, None]
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,S,W,idx", [(2, 4096, 32, 1, 3000), (2, 1000, 5, 3, 211), (3, 777, 16, 2, 5)])
def test_row_sharded_answer_combines_to_full_answer(world, n, S, W, idx):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, S, W, idx, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("rank hung")
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
