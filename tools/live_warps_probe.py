"""How does one CTA's time scale with the number of live tree warps?  A shard of k x 32 x 256 rows is one
depth-7 task with k live warps (the others publish empty slices).  Development aid."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
os.environ["PIRB_DFS_DEPTH"], os.environ["PIRB_HI_WARPS"] = "7", "16"
from pir_b200 import _capi
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
q0, q1 = _capi.PirQuery(), _capi.PirQuery()
tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
_capi.check(L.pir_client_gen_two_server(nb, 12345, 1, tape, C.byref(q0), C.byref(q1)))
print("FinalCW parity", q0.final_cw & 1)
for k in (16, 15, 14, 13, 12, 8, 4):
    cnt = k * 32 * 256
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, 0, cnt, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    a0 = np.zeros(S, np.uint8)
    for _ in range(3):
        _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
    _capi.check(L.pir_db_set_kernel_timing(h, 1))
    for _ in range(10):
        _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
    ms, c = C.c_double(), C.c_uint64()
    _capi.check(L.pir_db_get_kernel_timing(h, C.byref(ms), C.byref(c)))
    t = ms.value / c.value
    print("%2d live warps: %.4f ms  (%.4f ms per warp)" % (k, t, t / k), flush=True)
    L.pir_db_destroy(h)
