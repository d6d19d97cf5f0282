#!/bin/bash
# per-phase cycle counters of the answer kernel's tree warps (development aid)
export PIRB_PHASE_DEBUG=${PIRB_PHASE_DEBUG:-1}
python - <<'PY'
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from pir_b200 import _capi
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
rng = np.random.default_rng(1)
for frac, dmin in ((1, "4"), (1, "3"), (2, "3"), (8, "4"), (8, "2")):
    os.environ["PIRB_DMIN"] = dmin
    cnt = n // frac
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, 0, cnt, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
    _capi.check(L.pir_client_gen_two_server(nb, 12345, 1, tape, C.byref(q0), C.byref(q1)))
    a0 = np.zeros(S, np.uint8)
    print("shard 1/%d dmin %s" % (frac, dmin), file=sys.stderr, flush=True)
    for _ in range(3):
        _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), S))
    L.pir_db_destroy(h)
PY
