"""ctypes loader for oracle/libpir_oracle.so (TEST INFRASTRUCTURE, see oracle/__init__.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(ORACLE_DIR, "libpir_oracle.so")

PIR_MAX_NUM_BITS = 64
PIR_CW_BYTES = 18
TAPE_BYTES = 97


class PirQuery(C.Structure):
    """Mirror of `pir_query` in include/pir_b200.h (QueryShare + Key2P + PrfKeys flattened)."""

    _fields_ = [
        ("prf_keys", (C.c_uint8 * 16) * 4),
        ("s_init", C.c_uint8 * 16),
        ("t_init", C.c_uint8),
        ("share_number", C.c_uint8),
        ("is_keyword", C.c_uint8),
        ("is_two_party", C.c_uint8),
        ("num_bits", C.c_uint32),
        ("final_cw", C.c_int64),
        ("group_size", C.c_uint32),
        ("reserved", C.c_uint32),
        ("cw", (C.c_uint8 * PIR_CW_BYTES) * PIR_MAX_NUM_BITS),
    ]

    def copy(self):
        q = PirQuery()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(PirQuery))
        return q

    def to_bytes(self):
        return bytes(memoryview(self).cast("B"))


def query_from_bytes(b):
    q = PirQuery()
    assert len(b) == C.sizeof(PirQuery)
    C.memmove(C.byref(q), b, len(b))
    return q


def build(force=False):
    """Compile the C restatement with gcc (oracle/Makefile)."""
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("aes128.c", "pir_oracle.c", "aes128.h", "pir_oracle.h")]
    srcs.append(os.path.join(ORACLE_DIR, "..", "include", "pir_b200.h"))
    stale = force or not os.path.exists(_SO) or any(
        os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs if os.path.exists(s))
    if stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-B", "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_SO)
    u8p = C.POINTER(C.c_uint8)
    u64p = C.POINTER(C.c_uint64)
    qp = C.POINTER(PirQuery)
    L.orc_aes_selftest.restype = C.c_int
    L.orc_aes_using_aesni.restype = C.c_int
    L.orc_aes_force_portable.argtypes = [C.c_int]
    L.orc_aes128_expand_key.argtypes = [u8p, u8p]
    L.orc_aes128_encrypt.argtypes = [u8p, u8p, u8p]
    L.orc_varint8.argtypes = [u8p]
    L.orc_varint8.restype = C.c_int64
    L.orc_gen_two_server.argtypes = [u8p, C.c_uint32, C.c_uint64, C.c_uint64, qp, qp]
    L.orc_server_initialize.argtypes = [qp, C.c_uint32, C.c_void_p]
    L.orc_evaluate2p.argtypes = [C.c_void_p, qp, C.c_uint64]
    L.orc_evaluate2p.restype = C.c_int64
    L.orc_num_bits_for.argtypes = [C.c_uint64, C.c_uint32, C.c_int]
    L.orc_num_bits_for.restype = C.c_uint32
    L.orc_expand_shared_query.argtypes = [qp, C.c_uint64, u64p, C.c_uint64, C.c_uint64, u8p, C.c_int]
    L.orc_expand_tree.argtypes = [qp, C.c_uint64, C.c_uint64, C.c_uint64, u8p, C.c_int]
    L.orc_answer_with_bits.argtypes = [u8p, C.c_uint32, C.c_uint32, u8p, C.c_uint64, u8p]
    L.orc_answer.argtypes = [qp, u8p, C.c_uint64, C.c_uint32, u64p, u8p, C.c_int, C.c_int]
    L.orc_synth_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, u8p]
    L.orc_synth_word.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32]
    L.orc_synth_word.restype = C.c_uint64
    L.orc_answers_over_synth.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, u8p, C.c_uint32, u8p, C.c_int]
    L.orc_tape.argtypes = [C.c_uint64, u8p, C.c_size_t]
    if L.orc_aes_selftest() != 0:
        raise RuntimeError("oracle AES self-test failed (FIPS-197 vectors)")
    _lib = L
    return L


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _u64(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint64))


def tape(seed, n=TAPE_BYTES):
    out = np.zeros(n, dtype=np.uint8)
    lib().orc_tape(seed, _u8(out), n)
    return out


def num_bits_for(db_size, group_size=1, is_keyword=False):
    return int(lib().orc_num_bits_for(db_size, group_size, int(is_keyword)))


def gen_two_server(tape_bytes, num_bits, a, b=1):
    """dpf.ClientInitialize + GenerateTwoServer (dpf/client.go:17-150) on an explicit random tape."""
    t = np.ascontiguousarray(np.frombuffer(bytes(tape_bytes), dtype=np.uint8))
    assert t.size == TAPE_BYTES
    q0, q1 = PirQuery(), PirQuery()
    lib().orc_gen_two_server(_u8(t), num_bits, a, b, C.byref(q0), C.byref(q1))
    return q0, q1


def evaluate2p(q, xs, num_bits=None):
    """(*Dpf).Evaluate2P for each x in xs -> int64 array."""
    nb = q.num_bits if num_bits is None else num_bits
    f = C.create_string_buffer(3 * 176 + 8)
    lib().orc_server_initialize(C.byref(q), nb, f)
    xs = np.asarray(xs, dtype=np.uint64)
    out = np.empty(xs.size, dtype=np.int64)
    ev = lib().orc_evaluate2p
    for i, x in enumerate(xs.tolist()):
        out[i] = ev(f, C.byref(q), x)
    return out


def expand_shared_query(q, db_size, keywords=None, row_begin=0, n_rows=None, nthreads=1):
    """(*Database).ExpandSharedQuery (db.go:110-174): one byte per row, per-row Evaluate2P."""
    h = db_size // q.group_size
    if n_rows is None:
        n_rows = h - row_begin
    bits = np.zeros(max(n_rows, 1), dtype=np.uint8)
    kw = None
    if keywords is not None:
        kw = np.ascontiguousarray(keywords, dtype=np.uint64)
    lib().orc_expand_shared_query(C.byref(q), db_size, _u64(kw) if kw is not None else None,
                                  row_begin, n_rows, _u8(bits), nthreads)
    return bits[:n_rows]


def expand_tree(q, db_size, row_begin=0, n_rows=None, nthreads=1):
    h = db_size // q.group_size
    if n_rows is None:
        n_rows = h - row_begin
    bits = np.zeros(max(n_rows, 1), dtype=np.uint8)
    lib().orc_expand_tree(C.byref(q), db_size, row_begin, n_rows, _u8(bits), nthreads)
    return bits[:n_rows]


def answer_with_bits(slots, slot_bytes, group_size, bits):
    slots = np.ascontiguousarray(slots, dtype=np.uint8).reshape(-1)
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    out = np.zeros(group_size * slot_bytes, dtype=np.uint8)
    lib().orc_answer_with_bits(_u8(slots), slot_bytes, group_size, _u8(bits), bits.size, _u8(out))
    return out


def answer(q, slots, db_size, slot_bytes, keywords=None, nthreads=1, use_tree=False):
    """(*Database).PrivateSecretSharedQuery (db.go:68-72) over a dense host database."""
    slots = np.ascontiguousarray(slots, dtype=np.uint8).reshape(-1)
    assert slots.size >= db_size * slot_bytes
    out = np.zeros(q.group_size * slot_bytes, dtype=np.uint8)
    kw = None
    if keywords is not None:
        kw = np.ascontiguousarray(keywords, dtype=np.uint64)
    lib().orc_answer(C.byref(q), _u8(slots), db_size, slot_bytes, _u64(kw) if kw is not None else None,
                     _u8(out), nthreads, int(use_tree))
    return out


def answers_over_synth(seed, slot_bytes, group_size, bits_list, row_begin=0, nthreads=1):
    """db.go:75-107 for several bit vectors at once over the synthetic database `seed` (rows generated on the fly,
    nothing of the 4-8 GiB database is kept on the host).  bits_list: equal-length uint8 vectors for rows
    [row_begin, row_begin + len).  Returns one answer (group_size*slot_bytes bytes) per vector."""
    bits = np.ascontiguousarray(np.stack([np.asarray(b, dtype=np.uint8) for b in bits_list]))
    nq, n_rows = bits.shape
    out = np.zeros((nq, group_size * slot_bytes), dtype=np.uint8)
    lib().orc_answers_over_synth(seed, slot_bytes, group_size, row_begin, n_rows, _u8(bits), nq, _u8(out), nthreads)
    return out


def synth_fill(seed, slot_begin, n_slots, slot_bytes):
    out = np.zeros(n_slots * slot_bytes, dtype=np.uint8)
    lib().orc_synth_fill(seed, slot_begin, n_slots, slot_bytes, _u8(out))
    return out.reshape(n_slots, slot_bytes)
