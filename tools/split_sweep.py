#!/usr/bin/env python
"""Kernel time of the answer kernel for shard sizes 1, 1/2, 1/4, 1/8 of 2^24 x 256 B and every chunk depth Dmin
(PIRB_DMIN) plus the library's own choice; checks every answer against the XOR-recovered slot."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pir_b200 import _capi  # noqa: E402

L = _capi.lib()
n, S = 1 << int(os.environ.get("LOG_ROWS", "24")), int(os.environ.get("SLOT", "256"))
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
reps = 12
for frac in (1, 2, 4, 8):
    cnt = n // frac
    begin = (frac - 1) * cnt
    for dmin in ("auto", "1", "2", "3", "4", "5", "6"):
        if dmin == "auto":
            os.environ.pop("PIRB_DMIN", None)
        else:
            os.environ["PIRB_DMIN"] = dmin
        h = C.c_void_p()
        _capi.check(L.pir_db_create(0, n, S, begin, cnt, C.byref(h)))
        _capi.check(L.pir_db_fill_synthetic(h, 3))
        idx = begin + int(rng.integers(0, cnt))
        q0, q1 = _capi.PirQuery(), _capi.PirQuery()
        tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
        _capi.check(L.pir_client_gen_two_server(nb, idx, 1, tape, C.byref(q0), C.byref(q1)))
        a0, a1, want = np.zeros(S, np.uint8), np.zeros(S, np.uint8), np.zeros(S, np.uint8)
        _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
        _capi.check(L.pir_answer(h, C.byref(q1), a1.ctypes.data_as(C.c_void_p), S))
        _capi.check(L.pir_db_download(h, idx, want.ctypes.data_as(C.c_void_p), 1))
        ok = np.array_equal(a0 ^ a1, want)
        _capi.check(L.pir_db_set_kernel_timing(h, 1))
        for _ in range(reps):
            _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
        ms, cnt_l = C.c_double(), C.c_uint64()
        _capi.check(L.pir_db_get_kernel_timing(h, C.byref(ms), C.byref(cnt_l)))
        k = ms.value / cnt_l.value
        print("shard 1/%d dmin %-4s kernel %.4f ms  %.0f GB/s  ok=%s" % (frac, dmin, k, cnt * S / k / 1e6, ok), flush=True)
        L.pir_db_destroy(h)
