#!/usr/bin/env python
"""Back-to-back single-query launches on shards of 1, 1/2, 1/4, 1/8 of 2^24 x 256 B (what one GPU of an N-GPU server
runs), device-timed, for the four combinations of look-ahead and launch overlap (pir_db_set_option no_lookahead /
serial_launch).  Every answer of the default combination is checked against the XOR of the two servers' answers
recovering the slot.  Output: ms per query per combination -> profiles/r2_shard_probe.txt."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pir_b200 import _capi  # noqa: E402

if os.environ.get("VARIANT_SO"):     # A/B of library builds (tools/variant_shards.sh); a tool-side switch, not the product's
    _capi.LIB_PATH = os.environ["VARIANT_SO"]
L = _capi.lib()
n, S = 1 << int(os.environ.get("LOG_ROWS", "24")), int(os.environ.get("SLOT", "256"))
nb = L.pir_num_bits_for(n, 1, 0)
NQ = 40
rng = np.random.default_rng(7)
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)
for frac in (1, 2, 4, 8):
    cnt = n // frac
    begin = (frac - 1) * cnt
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, begin, cnt, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    qs = (_capi.PirQuery * (2 * NQ))()
    idxs = [int(x) for x in rng.integers(begin, begin + cnt, size=NQ)]
    for k, idx in enumerate(idxs):
        tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
        _capi.check(L.pir_client_gen_two_server(nb, idx, 1, tape, C.byref(qs[2 * k]), C.byref(qs[2 * k + 1])))
    d_out = torch.zeros((2 * NQ, S), dtype=torch.uint8, device="cuda")
    res = {}
    for la, ser in ((0, 0), (1, 0), (0, 1), (1, 1)):
        L.pir_db_set_option(h, b"no_lookahead", la)     # (a variant built before an option existed just ignores it)
        L.pir_db_set_option(h, b"serial_launch", ser)
        best = 1e9
        for rep in range(3):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for k in range(2 * NQ):
                _capi.check(L.pir_answer_batch_device(h, C.byref(qs[k]), 1, C.c_void_p(d_out[k].data_ptr()), sp))
            e1.record(st)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / (2 * NQ))
        res[(la, ser)] = best
        got = d_out.cpu().numpy()
        ok = True
        for k, idx in enumerate(idxs):
            want = np.zeros(S, np.uint8)
            _capi.check(L.pir_db_download(h, idx, want.ctypes.data_as(C.c_void_p), 1))
            ok = ok and np.array_equal(got[2 * k] ^ got[2 * k + 1], want)
        if not ok:
            print("shard 1/%d: WRONG ANSWER with no_lookahead=%d serial_launch=%d" % (frac, la, ser), flush=True)
    ideal = res[(1, 1)]
    print("shard 1/%d (%d rows): ms per query  default %.4f | no look-ahead %.4f | no overlap %.4f | neither %.4f   (%d keys x 2 servers back to back, best of 3)" % (
        frac, cnt, res[(0, 0)], res[(1, 0)], res[(0, 1)], res[(1, 1)], NQ), flush=True)
    L.pir_db_destroy(h)
