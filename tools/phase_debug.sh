#!/bin/bash
# per-phase cycle counters of the answer kernel's tree warps (development aid)
export PIRB_PHASE_DEBUG=${PIRB_PHASE_DEBUG:-1}
python - <<'PY'
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
from tools._keys import key_per_parity
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
cases = ((1, "b4"), (1, "7x12"), (2, "auto"), (8, "auto"))
if os.environ.get("PIRB_PHASE_CASES"):   # e.g. "8:auto,1:b4"
    cases = tuple((int(c.split(":")[0]), c.split(":")[1]) for c in os.environ["PIRB_PHASE_CASES"].split(","))
for frac, dmin in cases:
    for k in ("PIRB_DFS_DEPTH", "PIRB_HI_WARPS", "PIRB_DMIN"):
        os.environ.pop(k, None)
    if dmin.startswith("b"):
        os.environ["PIRB_DMIN"] = dmin[1:]
    elif dmin != "auto":
        os.environ["PIRB_DFS_DEPTH"], os.environ["PIRB_HI_WARPS"] = dmin.split("x")
    cnt = n // frac
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, 0, cnt, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    keys = key_per_parity(L, nb, 12345)
    a0 = np.zeros(S, np.uint8)
    for par in (0, 1):
        print("shard 1/%d shape %s FinalCW %s" % (frac, dmin, "odd" if par else "even"), file=sys.stderr, flush=True)
        for _ in range(2):
            _capi.check(L.pir_answer(h, C.byref(keys[par][0]), a0.ctypes.data_as(C.c_void_p), S))
    L.pir_db_destroy(h)
PY
