import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S = 1 << 24, 32
nb = L.pir_num_bits_for(n, 1, 0)
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
q0, q1 = _capi.PirQuery(), _capi.PirQuery()
tape = (C.c_uint8 * 97)(*range(97))
_capi.check(L.pir_client_gen_two_server(nb, 123456, 1, tape, C.byref(q0), C.byref(q1)))
bits = np.zeros(n, np.uint8)
res = {}
for mode in ("new", "v1"):
    if mode == "v1":
        os.environ["PIRB_EXPAND_V1"] = "1"
    else:
        os.environ.pop("PIRB_EXPAND_V1", None)
    ts = []
    for rep in range(4):
        t0 = time.perf_counter()
        _capi.check(L.pir_expand_bits(h, C.byref(q0), bits.ctypes.data_as(C.POINTER(C.c_uint8)), n))
        ts.append((time.perf_counter() - t0) * 1e3)
    res[mode] = bits.copy()
    print("pir_expand_bits 2^24 rows (%s): %.2f ms end to end (16 MiB of bools back to the host)" % (mode, min(ts)))
print("identical:", np.array_equal(res["new"], res["v1"]), "ones:", int(res["new"].sum()))
