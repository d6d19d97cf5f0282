import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
L = _capi.lib()
n, S, Q = 1 << 22, 1024, 64
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 3))
qs = (_capi.PirQuery * Q)(); scratch = _capi.PirQuery()
for k in range(Q):
    tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
    _capi.check(L.pir_client_gen_two_server(nb, int(rng.integers(0, n)), 1, tape, C.byref(qs[k]), C.byref(scratch)))
out = np.zeros(S * Q, np.uint8)
for rep in range(4):
    t0 = time.perf_counter()
    _capi.check(L.pir_answer_batch(h, qs, Q, out.ctypes.data_as(C.c_void_p), out.size))
    print("batch of %d: %.2f ms" % (Q, (time.perf_counter() - t0) * 1e3), flush=True)
for rep in range(3):
    t0 = time.perf_counter()
    for k in range(Q):
        _capi.check(L.pir_answer(h, C.byref(qs[k]), out.ctypes.data_as(C.c_void_p), S))
    print("64 single: %.2f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
