"""Seeded random sweep of database geometries through the C ABI against the oracle (bit-exact): heights that are not a
power of two, slots of any size, group sizes with a ragged tail (DBSize % GroupSize != 0, db.go:78-104's guard), shard
windows that start and end anywhere, index and keyword queries, both servers' shares, single calls and batches, and the
result-neutral path options of pir_db_set_option.  The reference's own tests draw their cases from crypto/rand
(db_test.go:18-64, dpf/dpf_test.go:11-107); here the draw is seeded so that a failure names a reproducible case."""
import numpy as np
import pytest

import oracle
from pir_b200 import sharding
from tests.helpers import Handle, make_queries

pytestmark = pytest.mark.gpu
CASES_PER_SEED = 5
OPTIONS = [None, None, ("kernel_v1", 1), ("expand_v1", 1), ("force_unfused", 1), ("keyword_v1", 1), ("no_lookahead", 1),
           ("serial_launch", 1), ("no_trim", 1), ("host_copy_path", 1), ("batch_mode", 1), ("batch_mode", 2)]


def _draw(rng):
    slot_bytes = int(rng.choice([int(rng.integers(1, 65)), int(rng.integers(65, 700)), 16 * int(rng.integers(1, 80)),
                                 int(rng.choice([32, 100, 256, 512, 1024, 1500]))]))
    db_size = int(2 ** rng.uniform(0, 17.5))
    db_size = max(1, min(db_size, (24 << 20) // slot_bytes))
    W = int(rng.choice([1, 1, 1, 2, 3, 4, 7]))
    W = min(W, db_size)
    rows = db_size // W
    if rng.integers(0, 3) == 0 or rows < 2:          # the whole database ...
        r0, cnt, to_end = 0, rows, True
    else:                                            # ... or a shard window of whole rows
        r0 = int(rng.integers(0, rows))
        cnt = int(rng.integers(1, rows - r0 + 1))
        to_end = r0 + cnt == rows
    slot_begin = r0 * W
    slot_count = db_size - slot_begin if to_end else cnt * W   # the last shard also holds the ragged tail
    return dict(db_size=db_size, slot_bytes=slot_bytes, W=W, rows=rows, r0=r0, cnt=cnt, slot_begin=slot_begin,
                slot_count=slot_count, keyword=bool(rng.integers(0, 4) == 0), opt=OPTIONS[int(rng.integers(0, len(OPTIONS)))],
                fill_seed=int(rng.integers(1, 1 << 30)), key_seed=int(rng.integers(1, 1 << 30)))


def _expected_part(data, c, bits):
    rows = data[:c["rows"] * c["W"]].reshape(c["rows"], c["W"] * c["slot_bytes"])[c["r0"]:c["r0"] + c["cnt"]]
    sel = rows[bits.astype(bool)]
    return np.bitwise_xor.reduce(sel, axis=0) if len(sel) else np.zeros(c["W"] * c["slot_bytes"], dtype=np.uint8)


@pytest.mark.parametrize("seed", range(40))
def test_random_geometries_match_the_oracle(gpu_lib, seed):
    rng = np.random.default_rng(1000 + seed)
    for k in range(CASES_PER_SEED):
        c = _draw(rng)
        tag = "seed %d case %d: %r" % (seed, k, c)
        data = oracle.synth_fill(c["fill_seed"], 0, c["db_size"], c["slot_bytes"])
        kws = None
        if c["keyword"]:
            kws = np.sort(rng.choice(1 << 32, size=c["db_size"], replace=False).astype(np.uint64))
            if rng.integers(0, 3) == 0:
                kws = kws[rng.permutation(c["db_size"])]                      # any order gives the same bits
        with Handle(c["db_size"], c["slot_bytes"], slot_begin=c["slot_begin"], slot_count=c["slot_count"]) as h:
            h.fill(c["fill_seed"])
            if c["opt"]:
                h.set_option(*c["opt"])
            if kws is not None:
                h.set_keywords(kws[c["slot_begin"]:c["slot_begin"] + c["slot_count"]])
                if c["W"] > 1 and c["r0"] > 0:
                    h.set_row_keywords(c["W"], kws[c["r0"]:c["r0"] + c["cnt"]])
            assert h.rows(c["W"]) == (c["r0"], c["cnt"]), tag
            # one target inside the window, one anywhere
            targets = [c["r0"] + int(rng.integers(0, c["cnt"])), int(rng.integers(0, c["rows"]))]
            for t, idx in enumerate(targets):
                a = int(kws[idx]) if kws is not None else idx
                (o0, o1), (q0, q1) = make_queries(c["key_seed"] + t, c["db_size"], c["W"], idx, is_keyword=kws is not None, keyword=a)
                parts = []
                for o, q in ((o0, q0), (o1, q1)):
                    want_bits = oracle.expand_shared_query(o, c["db_size"], keywords=kws, row_begin=c["r0"], n_rows=c["cnt"])
                    assert np.array_equal(h.expand_bits(q), want_bits), tag
                    want = _expected_part(data, c, want_bits)
                    got = h.answer(q)
                    assert np.array_equal(got, want), tag
                    assert np.array_equal(h.answer_with_bits(c["W"], want_bits), want), tag
                    parts.append(got)
                # Recover (query.go:278-297): the two shares' parts XOR to the target row if the window holds it, else to zero
                rec = parts[0] ^ parts[1]
                inside = c["r0"] <= idx < c["r0"] + c["cnt"]
                want_rec = data[idx * c["W"]:(idx + 1) * c["W"]].reshape(-1) if inside else np.zeros_like(rec)
                assert np.array_equal(rec, want_rec), tag
                both = h.answer_batch([q0, q1, q0, q1, q0])
                for j in range(5):
                    assert np.array_equal(both[j], parts[j & 1]), tag


@pytest.mark.parametrize("seed", range(12))
def test_random_row_sharded_servers_match_the_oracle(gpu_lib, seed):
    """The same draw through pir_server: 1-8 row shards (on as many GPUs as the box has, else sharing device 0), the in-kernel
    combine, index and keyword queries at the server's group size, single calls and batches."""
    rng = np.random.default_rng(5000 + seed)
    have = gpu_lib.pir_device_count()
    for k in range(4):
        c = _draw(rng)
        n_shards = int(rng.integers(1, 9))
        data = oracle.synth_fill(c["fill_seed"], 0, c["db_size"], c["slot_bytes"])
        kws = np.sort(rng.choice(1 << 32, size=c["db_size"], replace=False).astype(np.uint64)) if c["keyword"] else None
        with sharding.Server([i % have for i in range(n_shards)], c["db_size"], c["slot_bytes"], shard_align=c["W"]) as srv:
            srv.fill_synthetic(c["fill_seed"])
            opt = (OPTIONS + [("direct_combine", 1), ("signal_kernel", 1)])[int(rng.integers(0, len(OPTIONS) + 2))]
            if opt:
                srv.set_option(*opt)
            tag = "seed %d case %d: %d shards, server option %r, %r" % (seed, k, n_shards, opt, c)
            if kws is not None:
                srv.set_keywords(kws)
            answers = []
            for t in range(2):
                idx = int(rng.integers(0, c["rows"]))
                a = int(kws[idx]) if kws is not None else idx
                (o0, o1), (q0, q1) = make_queries(c["key_seed"] + t, c["db_size"], c["W"], idx, is_keyword=kws is not None, keyword=a)
                a0, a1 = srv.answer(q0), srv.answer(q1)
                assert np.array_equal(a0, oracle.answer(o0, data, c["db_size"], c["slot_bytes"], keywords=kws)), tag
                assert np.array_equal(a1, oracle.answer(o1, data, c["db_size"], c["slot_bytes"], keywords=kws)), tag
                assert np.array_equal(a0 ^ a1, data[idx * c["W"]:(idx + 1) * c["W"]].reshape(-1)), tag
                answers += [(q0, a0), (q1, a1)]
            both = srv.answer_batch([q for q, _ in answers])
            for j, (_, a) in enumerate(answers):
                assert np.array_equal(both[j], a), tag


@pytest.mark.parametrize("seed", range(24))
def test_random_medium_size_shards_match_the_oracle(gpu_lib, seed):
    """Sizes at which the answer kernel splits the tree over all CTAs in every way it knows (2^18 to 2^23.5 rows, up to 3 GiB):
    index queries on random shard windows, bits against the oracle's tree expansion and answers against its row XOR over the
    synthetic database, both shares, single calls and a batch."""
    rng = np.random.default_rng(9000 + seed)
    for k in range(3):
        slot_bytes = int(rng.choice([16, 32, 48, 100, 256, 336, 512, 1024]))
        W = int(rng.choice([1, 1, 2, 3]))
        rows = int(2 ** rng.uniform(18, 23.5))
        rows = min(rows, (3 << 30) // (slot_bytes * W))
        db_size = rows * W + int(rng.integers(0, W))                 # a ragged tail now and then
        r0 = int(rng.integers(0, rows // 2)) if rng.integers(0, 2) else 0
        cnt = int(rng.integers(rows // 4, rows - r0 + 1))
        to_end = r0 + cnt == rows
        opt = OPTIONS[int(rng.integers(0, len(OPTIONS)))]
        fill_seed, key_seed = int(rng.integers(1, 1 << 30)), int(rng.integers(1, 1 << 30))
        tag = "seed %d case %d: db_size %d slot_bytes %d W %d rows [%d, +%d) option %r" % (seed, k, db_size, slot_bytes, W, r0, cnt, opt)
        with Handle(db_size, slot_bytes, slot_begin=r0 * W, slot_count=db_size - r0 * W if to_end else cnt * W) as h:
            h.fill(fill_seed)
            if opt:
                h.set_option(*opt)
            assert h.rows(W) == (r0, cnt), tag
            idx = r0 + int(rng.integers(0, cnt))
            (o0, o1), (q0, q1) = make_queries(key_seed, db_size, W, idx)
            bits = [oracle.expand_tree(o, db_size, row_begin=r0, n_rows=cnt, nthreads=8) for o in (o0, o1)]
            want = oracle.answers_over_synth(fill_seed, slot_bytes, W, bits, row_begin=r0, nthreads=8)
            got = [h.answer(q0), h.answer(q1)]
            for s, q in enumerate((q0, q1)):
                assert np.array_equal(h.expand_bits(q), bits[s]), tag
                assert np.array_equal(got[s], want[s]), tag
            assert np.array_equal(got[0] ^ got[1], oracle.synth_fill(fill_seed, idx * W, W, slot_bytes).reshape(-1)), tag
            both = h.answer_batch([q1, q0, q0, q1])
            for j, s in enumerate((1, 0, 0, 1)):
                assert np.array_equal(both[j], want[s]), tag


@pytest.mark.parametrize("seed", range(6))
def test_arbitrary_key_bytes_evaluate_like_the_reference(gpu_lib, seed):
    """A share arrives from the network: whatever bytes it holds (PRF keys, seed, correction words with arbitrary tLCW / tRCW
    bytes, FinalCW), the server must compute what dpf/server.go:55-101 computes on them -- not only on keys an honest client
    minted.  Every field at random (flag fields within their ranges), index and keyword mode, against the oracle."""
    import ctypes as C
    from pir_b200 import _capi
    from tests.helpers import eval_points
    rng = np.random.default_rng(7000 + seed)
    for k in range(4):
        keyword = bool(k & 1)
        db_size, S, W = int(rng.integers(1, 40000)), int(rng.choice([16, 40, 256])), int(rng.choice([1, 2]))
        W = min(W, db_size)
        nb = oracle.num_bits_for(db_size, W, keyword)
        raw = bytearray(rng.integers(0, 256, size=C.sizeof(_capi.PirQuery), dtype=np.uint8).tobytes())
        q = _capi.PirQuery.from_buffer(raw)
        q.t_init, q.share_number, q.is_keyword, q.is_two_party = int(rng.integers(2)), int(rng.integers(2)), int(keyword), 1
        q.num_bits, q.group_size, q.reserved = nb, W, 0
        o = oracle.query_from_bytes(bytes(raw))
        tag = "seed %d case %d: db_size %d slot_bytes %d W %d keyword %r" % (seed, k, db_size, S, W, keyword)
        kws = rng.integers(0, 1 << 32, size=db_size, dtype=np.uint64) if keyword else None     # (duplicates allowed)
        xs = rng.integers(0, 1 << min(nb, 62), size=500, dtype=np.uint64)
        assert np.array_equal(eval_points(q, xs), oracle.evaluate2p(o, xs)), tag
        with Handle(db_size, S) as h:
            h.fill(9 + k)
            if keyword:
                h.set_keywords(kws)
            bits = oracle.expand_shared_query(o, db_size, keywords=kws)
            assert np.array_equal(h.expand_bits(q), bits), tag
            data = oracle.synth_fill(9 + k, 0, db_size, S)
            assert np.array_equal(h.answer(q), oracle.answer(o, data, db_size, S, keywords=kws)), tag
