"""Shared helpers for the parity tests: keys from the oracle's keygen, databases, C-ABI calls."""
import ctypes as C

import numpy as np

import oracle
from pir_b200 import _capi


def to_capi_query(oq, group_size=1, is_keyword=False):
    """oracle PirQuery -> product PirQuery (same POD bytes)."""
    q = _capi.PirQuery.from_bytes(oq.to_bytes())
    q.group_size = group_size
    q.is_keyword = 1 if is_keyword else 0
    return q


def make_queries(seed, db_size, group_size, index, is_keyword=False, keyword=None):
    """Both servers' queries for row `index` (or `keyword`), keys minted by the ORACLE keygen."""
    nb = oracle.num_bits_for(db_size, group_size, is_keyword)
    a = keyword if is_keyword else index
    o0, o1 = oracle.gen_two_server(oracle.tape(seed), nb, a, 1)
    for o in (o0, o1):
        o.group_size = group_size
        o.is_keyword = 1 if is_keyword else 0
    return (o0, o1), (to_capi_query(o0, group_size, is_keyword), to_capi_query(o1, group_size, is_keyword))


class Handle:
    def __init__(self, db_size, slot_bytes, slot_begin=0, slot_count=None, device=0):
        self.L = _capi.lib()
        self.db_size, self.slot_bytes = db_size, slot_bytes
        self.slot_begin = slot_begin
        self.slot_count = db_size - slot_begin if slot_count is None else slot_count
        self.h = C.c_void_p()
        _capi.check(self.L.pir_db_create(device, db_size, slot_bytes, slot_begin, self.slot_count, C.byref(self.h)))

    def upload(self, dense, slot_index=None):
        dense = np.ascontiguousarray(dense, dtype=np.uint8)
        n = dense.size // self.slot_bytes
        _capi.check(self.L.pir_db_upload(self.h, self.slot_begin if slot_index is None else slot_index,
                                         dense.ctypes.data_as(C.c_void_p), n))

    def download(self, slot_index, n):
        out = np.zeros((n, self.slot_bytes), dtype=np.uint8)
        _capi.check(self.L.pir_db_download(self.h, slot_index, out.ctypes.data_as(C.c_void_p), n))
        return out

    def fill(self, seed):
        _capi.check(self.L.pir_db_fill_synthetic(self.h, seed))

    def set_option(self, name, value):
        """pir_db_set_option: path selection / work-split override (results are bit-identical for every setting)."""
        _capi.check(self.L.pir_db_set_option(self.h, name.encode(), int(value)))

    def set_keywords(self, kw):
        kw = np.ascontiguousarray(kw, dtype=np.uint64)
        _capi.check(self.L.pir_db_set_keywords(self.h, kw.ctypes.data_as(C.POINTER(C.c_uint64)), kw.size))

    def set_row_keywords(self, group_size, kw):
        kw = np.ascontiguousarray(kw, dtype=np.uint64)
        _capi.check(self.L.pir_db_set_row_keywords(self.h, group_size, kw.ctypes.data_as(C.POINTER(C.c_uint64)), kw.size))

    def rows(self, group_size):
        a, b = C.c_uint64(), C.c_uint64()
        _capi.check(self.L.pir_db_rows(self.h, group_size, C.byref(a), C.byref(b)))
        return a.value, b.value

    def answer(self, q):
        n = q.group_size * self.slot_bytes
        out = np.zeros(n, dtype=np.uint8)
        _capi.check(self.L.pir_answer(self.h, C.byref(q), out.ctypes.data_as(C.c_void_p), n))
        return out

    def answer_batch(self, qs):
        n = qs[0].group_size * self.slot_bytes
        arr = (_capi.PirQuery * len(qs))(*qs)
        out = np.zeros(n * len(qs), dtype=np.uint8)
        _capi.check(self.L.pir_answer_batch(self.h, arr, len(qs), out.ctypes.data_as(C.c_void_p), out.size))
        return out.reshape(len(qs), n)

    def expand_bits(self, q):
        _, n = self.rows(q.group_size)
        bits = np.zeros(max(n, 1), dtype=np.uint8)
        _capi.check(self.L.pir_expand_bits(self.h, C.byref(q), bits.ctypes.data_as(C.POINTER(C.c_uint8)), n))
        return bits[:n]

    def answer_with_bits(self, group_size, bits):
        bits = np.ascontiguousarray(bits, dtype=np.uint8)
        n = group_size * self.slot_bytes
        out = np.zeros(n, dtype=np.uint8)
        _capi.check(self.L.pir_answer_with_bits(self.h, group_size, bits.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                bits.size, out.ctypes.data_as(C.c_void_p), n))
        return out

    def close(self):
        if self.h:
            self.L.pir_db_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def eval_points(q, xs, device=0):
    xs = np.ascontiguousarray(xs, dtype=np.uint64)
    out = np.zeros(xs.size, dtype=np.int64)
    _capi.check(_capi.lib().pir_dpf_eval_points(device, C.byref(q), xs.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                xs.size, out.ctypes.data_as(C.POINTER(C.c_int64))))
    return out
