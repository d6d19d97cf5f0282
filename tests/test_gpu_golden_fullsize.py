"""GPU parity against the committed golden vectors and, at the FULL sizes BASELINE.json names, against the oracle's
bytes (both shares): configs[2] 2^24 x 256 B, configs[3] 64 queries x 2^22 x 1 KiB, configs[4] 2^28 x 32 B index and
keyword — single handle and row-sharded pir_server.  The oracle side never holds the 4-8 GiB database: bits come from
oracle.expand_tree (db.go:110-174 by tree expansion, cross-checked against the per-row reference algorithm on
windows) and answers from oracle.answers_over_synth (db.go:75-107 over the seeded synthetic rows, generated on the
fly).  Semantics pinned: db.go:128-147 per row, db_test.go:18-64."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle
from pir_b200 import _capi, sharding
from tests.helpers import Handle, eval_points, make_queries

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "dpf_pir_vectors.json")))
NT = max(1, len(os.sched_getaffinity(0)))


# ---- the committed golden vectors, fed to the GPU directly ---------------------------------------------------------
@pytest.mark.parametrize("case", GOLDEN["pir"], ids=lambda c: c["name"])
def test_gpu_reproduces_golden_pir_case(gpu_lib, case):
    n, S, W = case["db_size"], case["slot_bytes"], case["group_size"]
    data = oracle.synth_fill(case["db_seed"], 0, n, S)
    rows = n // W
    for mode in ("upload", "fill"):
        with Handle(n, S) as h:
            if mode == "upload":
                h.upload(data)
            else:
                h.fill(case["db_seed"])
            answers = []
            for share in (0, 1):
                q = _capi.PirQuery.from_bytes(bytes.fromhex(case["query%d" % share]))
                want_bits = np.unpackbits(np.frombuffer(bytes.fromhex(case["bits%d_packed_le" % share]), dtype=np.uint8),
                                          bitorder="little")[:rows]
                want = bytes.fromhex(case["answer%d" % share])
                assert np.array_equal(h.expand_bits(q), want_bits)
                assert h.answer(q).tobytes() == want
                assert h.answer_with_bits(W, want_bits).tobytes() == want
                assert h.answer_batch([q, q, q])[2].tobytes() == want
                answers.append(np.frombuffer(want, dtype=np.uint8))
            idx = case["index"]
            assert np.array_equal(answers[0] ^ answers[1], data[idx * W:(idx + 1) * W].reshape(-1))
    if rows >= 2:  # the same case through a row-sharded server (2 shards on one GPU)
        with sharding.Server([0, 0], n, S, shard_align=W) as srv:
            srv.upload(data)
            for share in (0, 1):
                q = _capi.PirQuery.from_bytes(bytes.fromhex(case["query%d" % share]))
                assert srv.answer(q).tobytes() == bytes.fromhex(case["answer%d" % share])


@pytest.mark.parametrize("case", GOLDEN["dpf"], ids=lambda c: c["name"])
def test_gpu_reproduces_golden_dpf_case(gpu_lib, case):
    xs = np.array(case["points"], dtype=np.uint64)
    for share in (0, 1):
        q = _capi.PirQuery.from_bytes(bytes.fromhex(case["query%d" % share]))
        assert eval_points(q, xs).tolist() == case["values%d" % share]
        # one point per call, the shape of the reference's (*Dpf).Evaluate2P (dpf/server.go:55)
        assert [int(eval_points(q, xs[i:i + 1])[0]) for i in range(min(3, xs.size))] == case["values%d" % share][:3]


# ---- helpers -----------------------------------------------------------------------------------------------------
def _keys_of_both_parities(n, W=1, seed0=5000, want=(0, 1)):
    """Key pairs whose FinalCW is even / odd (the answer kernel has a code path per parity)."""
    out, seed = {}, seed0
    while len(out) < len(want):
        seed += 1
        idx = (seed * 2654435761) % (n // W)
        os_, qs = make_queries(seed, n, W, idx)
        par = qs[0].final_cw & 1
        if par in want and par not in out:
            out[par] = (idx, os_, qs)
    return [out[p] for p in want]


def _check_against_oracle(h, n, S, W, seed, idx, os_, qs, check_bits=True):
    """Both shares of one query: bits and answer equal the oracle's, and they recover the slot."""
    bits = [oracle.expand_tree(o, n, nthreads=NT) for o in os_]
    # the tree expansion agrees with the reference's per-row algorithm on windows (first, last, around the index)
    rows = n // W
    for r0 in (0, max(0, min(rows - 2048, idx - 1024)), rows - 2048):
        assert np.array_equal(bits[0][r0:r0 + 2048], oracle.expand_shared_query(os_[0], n, row_begin=r0, n_rows=2048))
    want = oracle.answers_over_synth(seed, S, W, bits, nthreads=NT)
    got = [h.answer(q) for q in qs]
    for share in (0, 1):
        if check_bits:
            assert np.array_equal(h.expand_bits(qs[share]), bits[share]), share
        assert np.array_equal(got[share], want[share]), share
    assert np.array_equal(got[0] ^ got[1], oracle.synth_fill(seed, idx * W, W, S).reshape(-1))
    return bits, want


# ---- configs[2]: 2^24 x 256 B ---------------------------------------------------------------------------------------
def test_config3_2pow24_x_256B_bit_exact_vs_oracle(gpu_lib):
    """The headline database at full size: ExpandSharedQuery bits and the answer bytes of both servers equal the
    oracle's, for a key of each FinalCW parity (both code paths run the shipped (7, 12) work split that only this
    size gets), and through pir_server with 2, 4 and 8 row shards."""
    n, S, seed = 1 << 24, 256, 2024
    cases = _keys_of_both_parities(n)
    wants = []
    with Handle(n, S) as h:
        h.fill(seed)
        for idx, os_, qs in cases:
            bits, want = _check_against_oracle(h, n, S, 1, seed, idx, os_, qs)
            wants.append(want)
            assert np.array_equal(h.answer_with_bits(1, bits[0]), want[0])
        many = h.answer_batch([c[2][0] for c in cases] * 3)
        for k in range(6):
            assert np.array_equal(many[k], wants[k % 2][0])
        # back-to-back single-query device calls on one stream: one launch each, consecutive launches overlap
        # (programmatic dependent launch) and share the handle's accumulator -- every answer must still be right,
        # also when the calls alternate between two streams
        import torch
        L = gpu_lib
        d_out = torch.zeros((24, S), dtype=torch.uint8, device="cuda")
        streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        for k in range(24):
            st = streams[0] if k < 12 else streams[k & 1]
            q = cases[(k // 2) & 1][2][k & 1]
            _capi.check(L.pir_answer_batch_device(h.h, C.byref(q), 1, C.c_void_p(d_out[k].data_ptr()), C.c_void_p(st.cuda_stream)))
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        for k in range(24):
            assert np.array_equal(got[k], wants[(k // 2) & 1][k & 1]), k
    for shards in (2, 4, 8):
        with sharding.Server([0] * shards, n, S) as srv:
            srv.fill_synthetic(seed)
            for (idx, os_, qs), want in zip(cases, wants):
                for share in (0, 1):
                    assert np.array_equal(srv.answer(qs[share]), want[share]), (shards, share)


# ---- configs[3]: 64 queries x 2^22 x 1 KiB ---------------------------------------------------------------------------
def test_config4_batch_of_64_over_2pow22_x_1KiB_vs_oracle(gpu_lib):
    n, S, seed, nq = 1 << 22, 1024, 404, 64
    rng = np.random.default_rng(64)
    idxs = [int(x) for x in rng.integers(0, n, size=nq)]
    pairs = [make_queries(7000 + i, n, 1, idx) for i, idx in enumerate(idxs)]
    os_ = [p[0][i & 1] for i, p in enumerate(pairs)]     # alternate shares
    qs = [p[1][i & 1] for i, p in enumerate(pairs)]
    bits = [oracle.expand_tree(o, n, nthreads=NT) for o in os_]
    want = np.concatenate([oracle.answers_over_synth(seed, S, 1, bits[k:k + 8], nthreads=NT) for k in range(0, nq, 8)])
    with Handle(n, S) as h:
        h.fill(seed)
        got = h.answer_batch(qs)                           # the two-phase path (bit vectors + one scan pass)
        assert np.array_equal(got, want)
        h.set_option("batch_mode", 1)                      # one fused launch per query
        assert np.array_equal(h.answer_batch(qs[:8]), want[:8])
        for k in range(0, nq, 7):
            assert np.array_equal(h.expand_bits(qs[k]), bits[k]), k
        # the other share of a few queries: recover
        for k in (0, 31, 63):
            other = h.answer(pairs[k][1][(k & 1) ^ 1])
            assert np.array_equal(other ^ got[k], oracle.synth_fill(seed, idxs[k], 1, S)[0])
    with sharding.Server([0, 0, 0, 0], n, S) as srv:
        srv.fill_synthetic(seed)
        assert np.array_equal(srv.answer_batch(qs[:16]), want[:16])


# ---- configs[4]: 2^28 x 32 B, index and keyword -----------------------------------------------------------------------
def _keywords_2p28(log_rows=28, seed=2024):
    """sorted, distinct 32-bit keywords: row i's keyword has i in its top bits and seeded low bits (bench.py's column)"""
    n, low = 1 << log_rows, 32 - log_rows
    out = np.empty(n, dtype=np.uint64)
    for c0 in range(0, n, 1 << 24):
        i = np.arange(c0, min(n, c0 + (1 << 24)), dtype=np.uint64)
        mix = (i * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed)) >> np.uint64(40)
        out[c0:c0 + i.size] = (i << np.uint64(low)) | (mix & np.uint64((1 << low) - 1))
    return out


def test_config5_2pow28_x_32B_index_and_keyword_vs_oracle(gpu_lib):
    n, S, seed = 1 << 28, 32, 28
    idx = 200_000_123
    os_, qs = make_queries(2800, n, 1, idx)
    kws = _keywords_2p28()
    kw_os, kw_qs = make_queries(2801, n, 1, idx, is_keyword=True, keyword=int(kws[idx]))
    windows = (0, idx - (1 << 19), n - (1 << 20))
    kw_want = [[oracle.expand_shared_query(o, n, keywords=kws, row_begin=r0, n_rows=1 << 20, nthreads=NT) for r0 in windows]
               for o in kw_os]
    with Handle(n, S) as h:
        h.fill(seed)
        bits, want = _check_against_oracle(h, n, S, 1, seed, idx, os_, qs)
        h.set_keywords(kws)
        kw_bits = [h.expand_bits(q) for q in kw_qs]
        for share in (0, 1):
            for r0, w in zip(windows, kw_want[share]):
                assert np.array_equal(kw_bits[share][r0:r0 + (1 << 20)], w), (share, r0)
        assert np.flatnonzero(kw_bits[0] ^ kw_bits[1]).tolist() == [idx]
        kw_ans = [h.answer(q) for q in kw_qs]
        # the answer over the GPU's own bits, folded by the oracle (the bits themselves are pinned on the windows above)
        kw_fold = oracle.answers_over_synth(seed, S, 1, kw_bits, nthreads=NT)
        assert np.array_equal(kw_ans[0], kw_fold[0]) and np.array_equal(kw_ans[1], kw_fold[1])
        assert np.array_equal(kw_ans[0] ^ kw_ans[1], oracle.synth_fill(seed, idx, 1, S)[0])
    del kw_bits, bits
    for shards in (2, 8):
        with sharding.Server([0] * shards, n, S) as srv:
            srv.fill_synthetic(seed)
            for share in (0, 1):
                assert np.array_equal(srv.answer(qs[share]), want[share]), (shards, share)
            if shards == 8:
                srv.set_keywords(kws)
                for share in (0, 1):
                    assert np.array_equal(srv.answer(kw_qs[share]), kw_ans[share]), share


# ---- error paths that must leave the handle usable ---------------------------------------------------------------------
@pytest.mark.parametrize("mode", [1, 2])
def test_bad_key_inside_a_batch_then_a_good_batch(gpu_lib, capi, mode):
    """A bad key at position >= 2 of a batch is rejected before anything is enqueued (round 1 returned from inside the
    launch loop with the earlier queries' kernels running on the side streams); the next batch is right."""
    n, S, nq = 1 << 13, 1024, 6
    data = oracle.synth_fill(31, 0, n, S)
    pairs = [make_queries(900 + i, n, 1, (i * 1237) % n) for i in range(nq)]
    qs = [p[1][0] for p in pairs]
    want = np.stack([oracle.answer(p[0][0], data, n, S) for p in pairs])
    with Handle(n, S) as h:
        h.fill(31)
        h.set_option("batch_mode", mode)
        for bad_at in (2, nq - 1):
            bad = list(qs)
            bad[bad_at] = _capi.PirQuery.from_bytes(qs[bad_at].to_bytes())
            bad[bad_at].t_init = 7
            for rep in range(3):
                with pytest.raises(capi.PirError) as e:
                    h.answer_batch(bad)
                assert e.value.code == capi.PIR_ERR_BAD_KEY
                assert np.array_equal(h.answer_batch(qs), want)


def test_server_error_at_a_later_shard_then_a_good_query(gpu_lib, capi):
    """A group size that shard 0 accepts and shard 1 does not: the call fails before any shard has launched, and the
    next query over the same server is right (pir_server_answer_batch returned from the shard loop)."""
    n, S = 1001, 48                      # 3 shards: [0, 333) [333, 667) [667, 1001); 9 divides 333, not 334
    data = oracle.synth_fill(17, 0, n, S)
    (o9, _), (q9, _) = make_queries(50, n, 9, 5)
    (o1, _), (q1, _) = make_queries(51, n, 1, 700)
    with sharding.Server([0, 0, 0], n, S, shard_align=1) as srv:
        srv.fill_synthetic(17)
        for rep in range(4):
            with pytest.raises(capi.PirError) as e:
                srv.answer(q9)
            assert e.value.code == capi.PIR_ERR_BAD_ARG
            assert np.array_equal(srv.answer(q1), oracle.answer(o1, data, n, S))
        # a bad key in the middle of a batch
        bad = _capi.PirQuery.from_bytes(q1.to_bytes())
        bad.share_number = 3
        with pytest.raises(capi.PirError):
            srv.answer_batch([q1, q1, bad, q1])
        assert np.array_equal(srv.answer_batch([q1, q1])[1], oracle.answer(o1, data, n, S))


def test_partly_uploaded_database_reads_zeros(gpu_lib):
    """Slots never uploaded are zeros (like make([]byte, n) in the reference), not stale device memory."""
    n, S = 4096, 256
    rng = np.random.default_rng(2)
    part = rng.integers(0, 256, size=(1000, S), dtype=np.uint8)
    data = np.zeros((n, S), dtype=np.uint8)
    data[500:1500] = part
    for rep in range(3):   # (a fresh handle each time re-uses freed device memory of the one before)
        with Handle(n, S) as h:
            h.upload(part, slot_index=500)
            (o0, _), (q0, _) = make_queries(60 + rep, n, 1, 777)
            assert np.array_equal(h.answer(q0), oracle.answer(o0, data, n, S))
            assert not h.download(2000, 16).any()
            h.fill(rep)    # leave non-zero bytes behind for the next handle


def test_options_reject_unknown_names_and_values(gpu_lib, capi):
    with Handle(64, 16) as h:
        with pytest.raises(capi.PirError):
            h.set_option("no_such_option", 1)
        with pytest.raises(capi.PirError):
            h.set_option("batch_mode", 9)
        h.set_option("batch_mode", 0)


def test_current_device_is_left_alone(gpu_lib):
    """No entry point changes the calling thread's current CUDA device."""
    L = gpu_lib
    L.pir_debug_current_device.restype = C.c_int
    before = L.pir_debug_current_device()
    devs = list(range(min(2, L.pir_device_count())))[::-1]     # the last device first
    with sharding.Server(devs, 4096, 32) as srv:
        srv.fill_synthetic(1)
        (_, _), (q0, _) = make_queries(5, 4096, 1, 9)
        srv.answer(q0)
        assert L.pir_debug_current_device() == before
        assert eval_points(q0, np.arange(4, dtype=np.uint64), device=devs[0]).size == 4
        assert L.pir_debug_current_device() == before
    comb = sharding.Combine(devs[0], 2, 64)
    comb.close()
    assert L.pir_debug_current_device() == before


def test_plain_c99_client_of_the_c_abi(gpu_lib, tmp_path):
    """tests/cabi_smoke.c: a C99 program (what cgo compiles against) linked to libpir_b200.so runs a golden case."""
    import subprocess
    case = next(c for c in GOLDEN["pir"] if c["name"] == "shared_query_1024x3_w3")
    n, S, W = case["db_size"], case["slot_bytes"], case["group_size"]
    data = oracle.synth_fill(case["db_seed"], 0, n, S)

    def arr(name, b):
        return "static const uint8_t %s[] = {%s};\n" % (name, ",".join(str(x) for x in bytes(b)))
    hdr = "#define G_NAME \"%s\"\n#define G_DB_SIZE %d\n#define G_SLOT_BYTES %d\n#define G_GROUP %d\n#define G_INDEX %d\n#define G_NUM_BITS %d\n" % (
        case["name"], n, S, W, case["index"], case["num_bits"])
    hdr += arr("G_TAPE", bytes.fromhex(case["tape"])) + arr("G_DB", data.tobytes())
    hdr += arr("G_ANSWER0", bytes.fromhex(case["answer0"])) + arr("G_ANSWER1", bytes.fromhex(case["answer1"]))
    hdr += arr("G_QUERY0", bytes.fromhex(case["query0"]))
    (tmp_path / "golden_case.h").write_text(hdr)
    exe = str(tmp_path / "cabi_smoke")
    libdir = os.path.dirname(_capi.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-I", str(tmp_path),
                           os.path.join(ROOT, "tests", "cabi_smoke.c"), "-o", exe, "-L", libdir, "-lpir_b200",
                           "-Wl,-rpath," + libdir])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "cabi_smoke ok" in r.stdout
