"""Kernel time of the answer kernel for keys with an even and an odd FinalCW (two code paths)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
if os.environ.get("VARIANT_SO"):     # A/B of library builds; a tool-side switch, not the product's
    _capi.LIB_PATH = os.environ["VARIANT_SO"]
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 3))
rng = np.random.default_rng(1)
found = {}
while len(found) < 2:
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
    _capi.check(L.pir_client_gen_two_server(nb, 1234567, 1, tape, C.byref(q0), C.byref(q1)))
    found.setdefault(q0.final_cw & 1, (q0, q1))
a = np.zeros(S, np.uint8); b = np.zeros(S, np.uint8); want = np.zeros(S, np.uint8)
_capi.check(L.pir_db_download(h, 1234567, want.ctypes.data_as(C.c_void_p), 1))
for name, val in [o.split("=") for o in os.environ.get("OPTS", "").split(",") if o]:   # e.g. OPTS=no_trim=1,no_lookahead=1
    _capi.check(L.pir_db_set_option(h, name.encode(), int(val)))
for par, (q0, q1) in sorted(found.items()):
    _capi.check(L.pir_answer(h, C.byref(q0), a.ctypes.data_as(C.c_void_p), S))
    _capi.check(L.pir_answer(h, C.byref(q1), b.ctypes.data_as(C.c_void_p), S))
    ok = np.array_equal(a ^ b, want)
    _capi.check(L.pir_db_set_kernel_timing(h, 1))
    for _ in range(10):
        _capi.check(L.pir_answer(h, C.byref(q0), a.ctypes.data_as(C.c_void_p), S))
    ms, cnt = C.c_double(), C.c_uint64()
    _capi.check(L.pir_db_get_kernel_timing(h, C.byref(ms), C.byref(cnt)))
    k = ms.value / cnt.value
    print("FinalCW %s: kernel %.4f ms, %.0f GB/s, recovered ok=%s" % ("odd" if par else "even", k, n * S / k / 1e6, ok), flush=True)
