#!/usr/bin/env python
"""Phase split of the two-phase batch (development aid): time the launches of a 64-query batch with the ncu-free
method of counting kernels by name is not available here, so time (a) bits-only by PIRB_BATCH_SKIP_SCAN and (b) all."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
if os.environ.get("DEV_SO"):
    _capi.DEV_LIB_PATH = os.path.abspath(os.environ["DEV_SO"])   # A/B of two builds
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S, Q = 1 << int(os.environ.get("LOG_ROWS", "22")), int(os.environ.get("SLOT", "1024")), int(os.environ.get("Q", "64"))
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
for mode in ("fused", "bits"):
    os.environ["PIRB_BATCH_MODE"] = mode
    for skip in ("", "1"):
        if mode == "fused" and skip:
            continue
        if skip:
            os.environ["PIRB_BATCH_SKIP_SCAN"] = "1"
        else:
            os.environ.pop("PIRB_BATCH_SKIP_SCAN", None)
        ts = []
        for rep in range(4):
            t0 = time.perf_counter()
            _capi.check(L.pir_answer_batch(h, qs, Q, out.ctypes.data_as(C.c_void_p), out.size))
            ts.append((time.perf_counter() - t0) * 1e3)
        print("mode %-5s skip_scan=%-1s  %.2f ms per batch of %d" % (mode, skip or "0", min(ts), Q), flush=True)
