import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
logn = int(os.environ.get("LOG_ROWS", "24"))
n, S = 1 << logn, 32
rng = np.random.default_rng(1)
# n distinct sorted 32-bit keywords, density n / 2^32
if logn >= 26:   # cheap generator for very large sets: one keyword in every window of 2^(32-logn)
    w = 32 - logn
    kws = (np.arange(n, dtype=np.uint64) << np.uint64(w)) + rng.integers(0, 1 << w, size=n, dtype=np.uint64)
else:
    kws = np.unique(rng.integers(0, 1 << 32, size=int(n * 1.01), dtype=np.uint64))[:n]
assert kws.size == n
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 3))
_capi.check(L.pir_db_set_keywords(h, kws.ctypes.data_as(C.POINTER(C.c_uint64)), n))
q0, q1 = _capi.PirQuery(), _capi.PirQuery()
tape = (C.c_uint8 * 97)(*range(97))
idx = 12345
_capi.check(L.pir_client_gen_two_server(32, int(kws[idx]), 1, tape, C.byref(q0), C.byref(q1)))
for q in (q0, q1):
    q.is_keyword = 1
a0, a1, want = np.zeros(S, np.uint8), np.zeros(S, np.uint8), np.zeros(S, np.uint8)
for mode in (("prefix-sharing",) if logn >= 26 else ("prefix-sharing", "v1")):
    if mode == "v1":
        os.environ["PIRB_KEYWORD_V1"] = "1"
    else:
        os.environ.pop("PIRB_KEYWORD_V1", None)
    ts = []
    for rep in range(3):
        t0 = time.perf_counter()
        _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
        ts.append((time.perf_counter() - t0) * 1e3)
    _capi.check(L.pir_answer(h, C.byref(q1), a1.ctypes.data_as(C.c_void_p), S))
    _capi.check(L.pir_db_download(h, idx, want.ctypes.data_as(C.c_void_p), 1))
    print("keyword query, 2^%d sorted keywords x %d B, %s: %.2f ms per query, recovered ok=%s" % (
        logn, S, mode, min(ts), np.array_equal(a0 ^ a1, want)), flush=True)
