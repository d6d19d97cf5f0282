"""Development helper: one key of each FinalCW parity (the answer kernel has a code path per parity)."""
import ctypes as C
import numpy as np
from pir_b200 import _capi


def key_per_parity(L, nb, index, seed=1):
    """{0: (q0, q1) with an even FinalCW, 1: (q0, q1) with an odd one} for the same queried index"""
    rng = np.random.default_rng(seed)
    found = {}
    while len(found) < 2:
        q0, q1 = _capi.PirQuery(), _capi.PirQuery()
        tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
        _capi.check(L.pir_client_gen_two_server(nb, index, 1, tape, C.byref(q0), C.byref(q1)))
        found.setdefault(q0.final_cw & 1, (q0, q1))
    return found
