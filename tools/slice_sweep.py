"""Kernel time of the answer kernel per shard size, key parity and slice size (PIRB_SLICE_LOG)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
from tools._keys import key_per_parity
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
keys = key_per_parity(L, nb, 12345)
out = np.zeros(S, np.uint8)
for frac in (1, 2, 4, 8):
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, 0, n // frac, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    for slog in ("2", "3"):
        os.environ["PIRB_SLICE_LOG"] = slog
        res = []
        for par in (0, 1):
            for _ in range(3):
                _capi.check(L.pir_answer(h, C.byref(keys[par][0]), out.ctypes.data_as(C.c_void_p), S))
            _capi.check(L.pir_db_set_kernel_timing(h, 1))
            for _ in range(20):
                _capi.check(L.pir_answer(h, C.byref(keys[par][0]), out.ctypes.data_as(C.c_void_p), S))
            ms, cnt = C.c_double(), C.c_uint64()
            _capi.check(L.pir_db_get_kernel_timing(h, C.byref(ms), C.byref(cnt)))
            _capi.check(L.pir_db_set_kernel_timing(h, 0))
            res.append(ms.value / cnt.value)
        print("shard 1/%d slice_log %s: even %.4f ms odd %.4f ms mean %.4f" % (frac, slog, res[0], res[1], (res[0] + res[1]) / 2), flush=True)
    L.pir_db_destroy(h)
