#!/usr/bin/env python
"""Kernel time of the answer kernel for shard sizes 1, 1/2, 1/4, 1/8 of 2^24 x 256 B and a set of task shapes
(uniform PIRB_DFS_DEPTH x PIRB_HI_WARPS, binary "bN" = PIRB_DMIN) plus the library's own choice; checks every answer against the XOR-recovered
slot.  SHAPES="auto 7x16 7x12 6x16" selects the shapes."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pir_b200 import _capi  # noqa: E402
from tools._keys import key_per_parity  # noqa: E402

_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only

L = _capi.lib()
n, S = 1 << int(os.environ.get("LOG_ROWS", "24")), int(os.environ.get("SLOT", "256"))
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
reps = 12
for frac in (1, 2, 4, 8):
    cnt = n // frac
    begin = (frac - 1) * cnt
    for dmin in os.environ.get("SHAPES", "auto 7x12 6x12 5x12 4x12 4x8 3x12 b5 b4 b3 b2 b1").split():
        for k in ("PIRB_DFS_DEPTH", "PIRB_HI_WARPS", "PIRB_DMIN"):
            os.environ.pop(k, None)
        if dmin.startswith("b"):    # binary split with chunks of 2^(10+Dmin) rows
            os.environ["PIRB_DMIN"] = dmin[1:]
        elif dmin != "auto":        # uniform split DxHI_WARPS
            os.environ["PIRB_DFS_DEPTH"], os.environ["PIRB_HI_WARPS"] = dmin.split("x")
        h = C.c_void_p()
        _capi.check(L.pir_db_create(0, n, S, begin, cnt, C.byref(h)))
        _capi.check(L.pir_db_fill_synthetic(h, 3))
        idx = begin + cnt // 3
        keys = key_per_parity(L, nb, idx)
        want = np.zeros(S, np.uint8)
        _capi.check(L.pir_db_download(h, idx, want.ctypes.data_as(C.c_void_p), 1))
        res = []
        for par in (0, 1):
            q0, q1 = keys[par]
            a0, a1 = np.zeros(S, np.uint8), np.zeros(S, np.uint8)
            _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
            _capi.check(L.pir_answer(h, C.byref(q1), a1.ctypes.data_as(C.c_void_p), S))
            ok = np.array_equal(a0 ^ a1, want)
            _capi.check(L.pir_db_set_kernel_timing(h, 1))
            for _ in range(reps):
                _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
            ms, cnt_l = C.c_double(), C.c_uint64()
            _capi.check(L.pir_db_get_kernel_timing(h, C.byref(ms), C.byref(cnt_l)))
            _capi.check(L.pir_db_set_kernel_timing(h, 0))
            res.append((ms.value / cnt_l.value, ok))
        mean = 0.5 * (res[0][0] + res[1][0])
        print("shard 1/%d shape %-5s kernel even %.4f odd %.4f mean %.4f ms  %.0f GB/s  ok=%s" % (
            frac, dmin, res[0][0], res[1][0], mean, cnt * S / mean / 1e6, res[0][1] and res[1][1]), flush=True)
        L.pir_db_destroy(h)
