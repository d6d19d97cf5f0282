"""Per-CTA phase counters of one answer launch (PIRB_PHASE_DEBUG=2) for a key of the given FinalCW parity."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
os.environ["PIRB_PHASE_DEBUG"] = "2"
from pir_b200 import _capi
from tools._keys import key_per_parity
_capi.use_dev_library()   # PIRB_* switches and phase counters exist in the -DPIRB_DEV build only
L = _capi.lib()
n, S = 1 << 24, 256
nb = L.pir_num_bits_for(n, 1, 0)
par = int(sys.argv[1]) if len(sys.argv) > 1 else 0
frac = int(sys.argv[2]) if len(sys.argv) > 2 else 1     # shard = the last 1/frac of the database
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, (frac - 1) * (n // frac), n // frac, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 3))
keys = key_per_parity(L, nb, 12345)
a0 = np.zeros(S, np.uint8)
for _ in range(2):
    print("---- launch", file=sys.stderr, flush=True)
    _capi.check(L.pir_answer(h, C.byref(keys[par][0]), a0.ctypes.data_as(C.c_void_p), S))
