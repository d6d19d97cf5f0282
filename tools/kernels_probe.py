"""One keyword query over 2^24 sorted keywords x 32 B and one 16-query two-phase batch over 2^20 x 1 KiB: the launches
that tools/capture_evidence.sh captures with ncu (keyword_walk_kernel, batch_scan_kernel); answers recover the slot."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pir_b200 import _capi  # noqa: E402

L = _capi.lib()
rng = np.random.default_rng(3)


def keys(nb, a, kw=False):
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
    _capi.check(L.pir_client_gen_two_server(nb, a, 1, tape, C.byref(q0), C.byref(q1)))
    q0.is_keyword = q1.is_keyword = 1 if kw else 0
    return q0, q1


# keyword query
n, S = 1 << 24, 32
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 5))
i = np.arange(n, dtype=np.uint64)
kws = np.ascontiguousarray((i << np.uint64(8)) | ((i * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(56)))
_capi.check(L.pir_db_set_keywords(h, kws.ctypes.data_as(C.POINTER(C.c_uint64)), n))
idx = 9_999_999
q0, q1 = keys(32, int(kws[idx]), True)
a0, a1, want = np.zeros(S, np.uint8), np.zeros(S, np.uint8), np.zeros(S, np.uint8)
_capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
_capi.check(L.pir_answer(h, C.byref(q1), a1.ctypes.data_as(C.c_void_p), S))
_capi.check(L.pir_db_download(h, idx, want.ctypes.data_as(C.c_void_p), 1))
print("keyword query over 2^24 x 32 B: recovered ok =", bool(np.array_equal(a0 ^ a1, want)), flush=True)
L.pir_db_destroy(h)
# two-phase batch
n, S, NQ = 1 << 20, 1024, 16
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 6))
nb = L.pir_num_bits_for(n, 1, 0)
idxs = [int(x) for x in rng.integers(0, n, size=NQ)]
pairs = [keys(nb, i) for i in idxs]
ok = True
outs = []
for share in (0, 1):
    arr = (_capi.PirQuery * NQ)(*[p[share] for p in pairs])
    out = np.zeros((NQ, S), np.uint8)
    _capi.check(L.pir_answer_batch(h, arr, NQ, out.ctypes.data_as(C.c_void_p), out.size))
    outs.append(out)
for k, i in enumerate(idxs):
    w = np.zeros(S, np.uint8)
    _capi.check(L.pir_db_download(h, i, w.ctypes.data_as(C.c_void_p), 1))
    ok = ok and np.array_equal(outs[0][k] ^ outs[1][k], w)
print("16-query two-phase batch over 2^20 x 1 KiB: recovered ok =", bool(ok), flush=True)
L.pir_db_destroy(h)
