"""CPU tests of the product's host side: the C-ABI library loads and exports every symbol of
include/pir_b200.h, host-only entry points agree with the oracle / golden vectors, and compute entry
points FAIL LOUDLY without a CUDA device (there is no CPU fallback).  No kernel is launched here."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import oracle
from pir_b200 import _capi, sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "dpf_pir_vectors.json")))


def test_header_symbols_all_exported():
    hdr = open(os.path.join(ROOT, "include", "pir_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pir_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    L = _capi.lib()
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert declared == set(_capi.EXPORTS), declared ^ set(_capi.EXPORTS)
    assert L.pir_abi_version() == 1


def test_pod_layout_matches_oracle_pod():
    assert C.sizeof(_capi.PirQuery) == C.sizeof(oracle.PirQuery) == 64 + 16 + 4 + 4 + 8 + 4 + 4 + 64 * 18
    for (n1, _), (n2, _) in zip(_capi.PirQuery._fields_, oracle.PirQuery._fields_):
        assert n1 == n2
        assert getattr(_capi.PirQuery, n1).offset == getattr(oracle.PirQuery, n2).offset


@pytest.mark.parametrize("case", GOLDEN["pir"] + GOLDEN["dpf"], ids=lambda c: c["name"])
def test_client_keygen_matches_golden(case):
    """pir_client_gen_two_server (dpf/client.go:17-150) reproduces the golden keys from the same tape."""
    L = _capi.lib()
    tape = (C.c_uint8 * 97).from_buffer_copy(bytes.fromhex(case["tape"]))
    a = case.get("index", case.get("a"))
    b = case.get("b", 1)
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    _capi.check(L.pir_client_gen_two_server(case["num_bits"], a, b, tape, C.byref(q0), C.byref(q1)))
    for q, want in ((q0, case["query0"]), (q1, case["query1"])):
        q.group_size = case.get("group_size", 1)
        assert q.to_bytes().hex() == want


def test_client_keygen_random_tape_roundtrips_through_oracle():
    L = _capi.lib()
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    _capi.check(L.pir_client_gen_two_server(12, 3000, 1, None, C.byref(q0), C.byref(q1)))
    xs = np.arange(4096, dtype=np.uint64)
    s = oracle.evaluate2p(oracle.query_from_bytes(q0.to_bytes()), xs) + \
        oracle.evaluate2p(oracle.query_from_bytes(q1.to_bytes()), xs)
    assert s[3000] == 1 and np.count_nonzero(s) == 1


def test_num_bits_for_matches_reference_rule():
    L = _capi.lib()
    for n in (1, 2, 3, 4, 1023, 1024, 1025, 1 << 20, (1 << 24) - 1, 1 << 24, 1 << 28):
        for W in (1, 2, 3, 7):
            if n // W == 0:
                continue
            assert L.pir_num_bits_for(n, W, 0) == (n // W).bit_length() == oracle.num_bits_for(n, W, False)
            assert L.pir_num_bits_for(n, W, 1) == 32


def test_round_keys_and_ttable_vs_fips197():
    L = _capi.lib()
    key = (C.c_uint8 * 16)(*bytes.fromhex("000102030405060708090a0b0c0d0e0f"))
    rk = (C.c_uint32 * 44)()
    L.pir_debug_round_keys(key, rk)
    assert rk[43].to_bytes(4, "little").hex() == "4d2b30c5"  # FIPS-197 C.1 round[10].k_sch = 13111d7fe3944a17f307a78b4d2b30c5
    orc = np.zeros(176, dtype=np.uint8)
    oracle.lib().orc_aes128_expand_key(key, orc.ctypes.data_as(C.POINTER(C.c_uint8)))
    assert np.array_equal(np.frombuffer(bytes(rk), dtype=np.uint8), orc)
    t0 = (C.c_uint32 * 256)()
    L.pir_debug_t0(t0)
    assert t0[0] == 0xA56363C6 and t0[1] == 0x847C7CF8  # S(0)=0x63: [2s, s, s, 3s] little-endian


def test_xor_combine():
    rng = np.random.default_rng(0)
    parts = [rng.integers(0, 256, size=777, dtype=np.uint8) for _ in range(5)]
    want = parts[0] ^ parts[1] ^ parts[2] ^ parts[3] ^ parts[4]
    assert np.array_equal(sharding.xor_combine(parts), want)
    with pytest.raises(ValueError):
        sharding.xor_combine([b"ab", b"abc"])


def test_shard_slots_cover_database_exactly():
    for n, W in ((1 << 24, 1), (1000, 3), (1024, 4), (7, 2), (5, 8), (12345, 64)):
        for world in (1, 2, 3, 4, 8):
            pos, rows = 0, 0
            for r in range(world):
                b, c = sharding.shard_slots(n, W, r, world)
                assert b == pos and b % W == 0
                assert c % W == 0 or r == world - 1
                pos += c
                rows += sharding.shard_rows(n, W, r, world)[1]
            assert pos == n and rows == n // W


@pytest.mark.skipif(_capi.lib().pir_device_count() > 0, reason="a CUDA device is present")
def test_compute_entry_points_fail_loudly_without_gpu():
    L = _capi.lib()
    h = C.c_void_p()
    rc = L.pir_db_create(0, 1024, 16, 0, 1024, C.byref(h))
    assert rc == _capi.PIR_ERR_CUDA and not h.value
    assert b"no CUDA device" in L.pir_last_error()
    q = _capi.PirQuery.from_bytes(bytes.fromhex(GOLDEN["dpf"][1]["query0"]))
    xs = (C.c_uint64 * 2)(0, 1)
    out = (C.c_int64 * 2)()
    assert L.pir_dpf_eval_points(0, C.byref(q), xs, 2, out) == _capi.PIR_ERR_CUDA
    with pytest.raises(_capi.PirError):
        _capi.check(L.pir_db_create(0, 1024, 16, 0, 1024, C.byref(h)))


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under pir_b200/ may reference it."""
    pkg = os.path.join(ROOT, "pir_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")) or f == "Makefile":
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "pir_oracle" not in src and "orc_" not in src, f


def test_auth_token_shares_xor_to_key_and_check_audit():
    """aspir.go:227-243, 286-299 (host arithmetic, no device)."""
    from pir_b200 import aspir, pir
    key = pir.NewRandomSlot(16)
    for n in (2, 3, 5):
        shares = aspir.NewAuthTokenSharesForKey(key, n)
        acc = pir.NewEmptySlot(16)
        for s in shares:
            pir.XorSlots(acc, s.T)
        assert acc.Equal(key)
    a, b = pir.NewRandomSlot(16), pir.NewRandomSlot(16)
    assert aspir.CheckAudit(aspir.AuditTokenShare(a), aspir.AuditTokenShare(pir.Slot(bytes(a.Data))))
    assert not aspir.CheckAudit(aspir.AuditTokenShare(a), aspir.AuditTokenShare(b))


def test_multi_gpu_entry_points_fail_loudly_without_device():
    L = _capi.lib()
    if L.pir_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    h = C.c_void_p()
    devs = (C.c_int * 2)(0, 1)
    assert L.pir_server_create(devs, 2, 1024, 32, 1, C.byref(h)) == _capi.PIR_ERR_CUDA
    assert b"no CUDA device" in L.pir_last_error()
    assert L.pir_combine_create(0, 4, 256, C.byref(h)) == _capi.PIR_ERR_CUDA
    # argument validation happens before any device work
    assert L.pir_combine_create(0, 0, 256, C.byref(h)) == _capi.PIR_ERR_BAD_ARG
    assert L.pir_combine_create(0, 4, 100, C.byref(h)) == _capi.PIR_ERR_BAD_ARG
    assert L.pir_server_create(devs, 0, 1024, 32, 1, C.byref(h)) == _capi.PIR_ERR_BAD_ARG


def test_work_split_chooser_is_valid_for_typical_shards():
    """choose_answer_split (capi.cu) for the shards of the BASELINE configs: the split it returns is one the answer
    kernel accepts (launch_answer's checks), and the headline database gets one task per CTA at every shard size."""
    L = _capi.lib()
    f = L.pir_debug_answer_split
    f.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64, C.c_int] + [C.POINTER(C.c_uint32)] * 3
    f.restype = None
    sms = 148
    for logn in (10, 16, 20, 22, 24, 26, 28):
        for frac in (1, 2, 4, 8):
            n = (1 << logn) // frac
            d, dmin, hiw = C.c_uint32(), C.c_uint32(), C.c_uint32()
            f(logn + 1, (frac - 1) * n, n, sms, C.byref(d), C.byref(dmin), C.byref(hiw))
            assert d.value <= 7 and d.value + 1 <= logn + 1
            assert hiw.value in (0, 4, 8, 12, 16)
            if hiw.value == 0:
                assert dmin.value <= d.value
            else:
                assert hiw.value == 16 or d.value > 0
                if logn == 24:  # one uniform task per CTA: R = 2^(D+5) * (16 + hi_warps) rows cover the shard in <= 148 tasks
                    rows_per_task = (16 + hiw.value) << (d.value + 5)
                    assert (n + rows_per_task - 1) // rows_per_task <= sms
            if logn == 24:
                assert hiw.value == 12 and d.value == 7 - (frac.bit_length() - 1)


# ---- wire encoding of a query share (include/pir_b200.h: pir_query_wire_size / _encode / _decode) ----------------
def _encode(L, q):
    buf = (C.c_uint8 * L.pir_query_wire_size(q.num_bits))()
    n = C.c_size_t()
    _capi.check(L.pir_query_encode(C.byref(q), buf, len(buf), C.byref(n)))
    assert n.value == len(buf) == 100 + 18 * q.num_bits
    return bytes(buf)


def _decode(L, data):
    q = _capi.PirQuery()
    raw = (C.c_uint8 * max(len(data), 1)).from_buffer_copy(bytes(data) or b"\0")
    _capi.check(L.pir_query_decode(raw, len(data), C.byref(q)))
    return q


@pytest.mark.parametrize("case", GOLDEN["pir"] + GOLDEN["dpf"], ids=lambda c: c["name"])
def test_wire_encoding_round_trips_every_golden_share(case):
    L = _capi.lib()
    for key in ("query0", "query1"):
        q = _capi.PirQuery.from_buffer_copy(bytes.fromhex(case[key]))
        wire = _encode(L, q)
        # the layout is the documented one, byte for byte, written here independently of the C code
        want = b"PIRQ" + bytes([1, q.share_number | q.is_keyword << 1 | q.t_init << 2 | q.is_two_party << 3, q.num_bits, 0])
        want += int(q.group_size).to_bytes(4, "little") + int(q.final_cw).to_bytes(8, "little", signed=True)
        want += bytes(q.s_init) + b"".join(bytes(q.prf_keys[i]) for i in range(4))
        want += b"".join(bytes(q.cw[i]) for i in range(q.num_bits))
        assert wire == want
        assert _decode(L, wire).to_bytes() == q.to_bytes()


def test_wire_encoding_covers_every_field_and_size():
    L = _capi.lib()
    rng = np.random.default_rng(77)
    assert L.pir_query_wire_size(0) == 0 and L.pir_query_wire_size(65) == 0
    for nb in (1, 17, 25, 29, 32, 64):
        q = _capi.PirQuery()
        C.memmove(C.byref(q), rng.integers(0, 256, size=C.sizeof(q), dtype=np.uint8).tobytes(), C.sizeof(q))
        q.t_init, q.share_number, q.is_keyword, q.is_two_party = int(rng.integers(2)), int(rng.integers(2)), int(rng.integers(2)), 1
        q.num_bits, q.reserved = nb, 0
        q.final_cw = int(rng.integers(-2**63, 2**63 - 1))
        for i in range(nb, 64):                      # correction words beyond num_bits are not part of the share
            C.memset(q.cw[i], 0, 18)
        assert L.pir_query_wire_size(nb) == 100 + 18 * nb
        assert _decode(L, _encode(L, q)).to_bytes() == q.to_bytes()


def test_wire_decoding_rejects_malformed_input():
    L = _capi.lib()
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    _capi.check(L.pir_client_gen_two_server(17, 41000, 1, None, C.byref(q0), C.byref(q1)))
    good = _encode(L, q1)
    bad = {
        "empty": b"", "short header": good[:99], "truncated": good[:-1], "trailing byte": good + b"\0",
        "magic": b"QRIP" + good[4:], "version": good[:4] + b"\x02" + good[5:], "flag padding": good[:5] + bytes([good[5] | 0x10]) + good[6:],
        "pad byte": good[:7] + b"\x01" + good[8:], "num_bits 0": good[:6] + b"\x00" + good[7:], "num_bits 65": good[:6] + b"\x41" + good[7:],
        "num_bits vs length": good[:6] + b"\x12" + good[7:],
    }
    for name, data in bad.items():
        q = _capi.PirQuery()
        C.memset(C.byref(q), 0xEE, C.sizeof(q))
        raw = (C.c_uint8 * max(len(data), 1)).from_buffer_copy(data or b"\0")
        assert L.pir_query_decode(raw, len(data), C.byref(q)) == _capi.PIR_ERR_BAD_KEY, name
        assert L.pir_last_error()
        assert q.to_bytes() == bytes(C.sizeof(q)), name        # nothing half-decoded is left behind
    # encode: a buffer one byte short, flag fields out of range
    buf = (C.c_uint8 * (len(good) - 1))()
    n = C.c_size_t(5)
    assert L.pir_query_encode(C.byref(q1), buf, len(buf), C.byref(n)) == _capi.PIR_ERR_BAD_ARG and n.value == 0
    q1.t_init = 2
    buf = (C.c_uint8 * len(good))()
    assert L.pir_query_encode(C.byref(q1), buf, len(buf), None) == _capi.PIR_ERR_BAD_KEY


def test_query_share_marshal_binary_mirror():
    """pir.QueryShare.MarshalBinary / UnmarshalBinary: the share a server decodes evaluates like the one the client made."""
    from pir_b200 import pir
    md = pir.DBMetadata(SlotBytes=32, DBSize=3000)
    for shares in (md.NewIndexQueryShares(1234, 2, 2), md.NewKeywordQueryShares(99, 1, 2)):
        for s in shares:
            wire = s.MarshalBinary()
            back = pir.QueryShare.UnmarshalBinary(wire)
            assert back.to_pod().to_bytes() == s.to_pod().to_bytes()
            assert (back.GroupSize, back.IsKeywordBased, back.ShareNumber) == (s.GroupSize, s.IsKeywordBased, s.ShareNumber)
            assert back.MarshalBinary() == wire
    with pytest.raises(_capi.PirError):
        pir.QueryShare.UnmarshalBinary(b"not a share")
