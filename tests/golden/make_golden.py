#!/usr/bin/env python
"""Generate tests/golden/dpf_pir_vectors.json.

Run from the repo root:  python tests/golden/make_golden.py

The reference (Go) cannot run in the build container (no Go toolchain) and holds no known-answer
vectors for this path, so the vectors are produced by TWO independent restatements
of the cited Go lines and are only written when both agree byte-for-byte:
  * oracle/pir_oracle.c      (C, own AES-128 / AES-NI)
  * oracle/py_restatement.py (pure Python, AES from the `cryptography` package / OpenSSL)
Each case fixes the 97 bytes the reference would draw from crypto/rand (dpf/client.go:17-76).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import py_restatement as pr  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dpf_pir_vectors.json")

# (name, tape_seed, db_size, slot_bytes, group_size, index)
PIR_CASES = [
    ("shared_query_1024x3_w1", 101, 1024, 3, 1, 517),     # db_test.go:18-64 shapes
    ("shared_query_1024x3_w2", 102, 1024, 3, 2, 300),
    ("shared_query_1024x3_w3", 103, 1024, 3, 3, 340),     # ragged: 1024 % 3 != 0
    ("shared_query_1024x3_w4", 104, 1024, 3, 4, 255),
    ("one_row", 105, 1, 16, 1, 0),
    ("odd_height_777x32", 106, 777, 32, 1, 776),
    ("aspir_keydb_512x8", 107, 512, 8, 1, 77),            # aspir_test.go:62-64 slot size
    ("wide_slots_300x256", 108, 300, 256, 1, 299),
    ("group_20x100_w7", 109, 140, 100, 7, 19),
]
# (name, tape_seed, num_bits, a, b, points)
DPF_CASES = [
    ("dpf_nb1", 201, 1, 1, 1, [0, 1]),
    ("dpf_nb7", 202, 7, 99, 1, list(range(0, 128, 5)) + [99]),
    ("dpf_nb11_b5", 203, 11, 1123, 5, [0, 1, 1122, 1123, 1124, 2047]),
    ("dpf_nb32_keyword", 204, 32, 0xDEADBEEF, 1, [0, 1, 0xDEADBEEF, 0xDEADBEEE, 0xFFFFFFFF, 0x80000000]),
    ("dpf_nb64", 205, 64, 0x0123456789ABCDEF, 1, [0, 0x0123456789ABCDEF, 0xFFFFFFFFFFFFFFFF, 1 << 63]),
]


def py_key(k):
    return k


def check_keys(oq, keys, k):
    """C-oracle pir_query vs Python key dict."""
    assert bytes(bytearray(oq.s_init)) == k["SInit"]
    assert oq.t_init == k["TInit"]
    assert oq.final_cw == k["FinalCW"]
    for i, cw in enumerate(k["CW"]):
        assert bytes(bytearray(oq.cw[i])) == cw, i
    for i in range(4):
        assert bytes(bytearray(oq.prf_keys[i])) == keys[i]


def pir_case(name, seed, db_size, slot_bytes, W, index):
    tape = oracle.tape(seed)
    nb = oracle.num_bits_for(db_size, W, False)
    o0, o1 = oracle.gen_two_server(tape, nb, index, 1)
    keys, k0, k1 = pr.generate_two_server(tape, nb, index, 1)
    check_keys(o0, keys, k0)
    check_keys(o1, keys, k1)
    data = oracle.synth_fill(seed, 0, db_size, slot_bytes)
    slots = [bytes(data[i]) for i in range(db_size)]
    shares, bitsets = [], []
    for share, (o, k) in enumerate(((o0, k0), (o1, k1))):
        o.group_size = W
        bits_c = oracle.expand_shared_query(o, db_size)
        bits_t = oracle.expand_tree(o, db_size)
        bits_p = pr.expand_shared_query(keys, k, share, db_size, W)
        assert bits_c.tolist() == [int(b) for b in bits_p], name
        assert np.array_equal(bits_c, bits_t), name
        ans_c = oracle.answer(o, data, db_size, slot_bytes)
        ans_p = pr.answer_with_bits(slots, slot_bytes, W, bits_p)
        assert bytes(ans_c) == ans_p, name
        shares.append(ans_p.hex())
        bitsets.append(np.packbits(bits_c, bitorder="little").tobytes().hex())
    rec = bytes(a ^ b for a, b in zip(bytes.fromhex(shares[0]), bytes.fromhex(shares[1])))
    want = b"".join(slots[index * W + c] for c in range(W))
    assert rec == want, name
    return {"name": name, "tape": bytes(tape).hex(), "db_seed": seed, "db_size": db_size,
            "slot_bytes": slot_bytes, "group_size": W, "index": index, "num_bits": nb,
            "query0": o0.to_bytes().hex(), "query1": o1.to_bytes().hex(),
            "bits0_packed_le": bitsets[0], "bits1_packed_le": bitsets[1],
            "answer0": shares[0], "answer1": shares[1]}


def dpf_case(name, seed, nb, a, b, points):
    tape = oracle.tape(seed)
    o0, o1 = oracle.gen_two_server(tape, nb, a, b)
    keys, k0, k1 = pr.generate_two_server(tape, nb, a, b)
    check_keys(o0, keys, k0)
    check_keys(o1, keys, k1)
    prf = pr.Prf(keys)
    vals = []
    for share, (o, k) in enumerate(((o0, k0), (o1, k1))):
        vc = oracle.evaluate2p(o, np.array(points, dtype=np.uint64)).tolist()
        vp = [pr.evaluate_2p(prf, nb, share, k, x) for x in points]
        assert vc == vp, name
        vals.append(vp)
    for x, v0, v1 in zip(points, *vals):
        assert pr._wrap(v0 + v1) == (b if x == a else 0), (name, x)
    return {"name": name, "tape": bytes(tape).hex(), "num_bits": nb, "a": a, "b": b, "points": points,
            "query0": o0.to_bytes().hex(), "query1": o1.to_bytes().hex(),
            "values0": vals[0], "values1": vals[1]}


def main():
    doc = {
        "about": "two-restatement-agreed vectors for the DPF-PIR answer path; see make_golden.py",
        "aes_fips197_c1": {"key": "000102030405060708090a0b0c0d0e0f", "pt": "00112233445566778899aabbccddeeff",
                           "ct": "69c4e0d86a7b0430d8cdb78070b4c55a"},
        "xor_slots": [  # slot_test.go:146-175, the reference's only literal vectors
            {"a": "00000000", "b": "01010101", "out": "01010101"},
            {"a": "00000000", "b": "01000100", "out": "01000100"},
        ],
        "pir": [pir_case(*c) for c in PIR_CASES],
        "dpf": [dpf_case(*c) for c in DPF_CASES],
    }
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
