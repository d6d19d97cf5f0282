#!/usr/bin/env python
"""Kernel time per query for row sizes that are not a power of two (2^22 rows): the warp-specialised answer kernel
(shipped) against the round-1 two-launch path (pir_db_set_option kernel_v1).  Answers of both must agree."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pir_b200 import _capi  # noqa: E402

L = _capi.lib()
rng = np.random.default_rng(5)
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)
for slot_bytes, W in ((100, 3), (48, 1), (100, 1), (256, 1), (1008, 1)):
    rows = 1 << 22
    n = rows * W
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, slot_bytes, 0, n, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 3))
    pitch = L.pir_db_slot_pitch(h)
    nb = L.pir_num_bits_for(n, W, 0)
    qs = (_capi.PirQuery * 16)()
    scratch = _capi.PirQuery()
    for k in range(16):
        tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
        _capi.check(L.pir_client_gen_two_server(nb, int(rng.integers(0, rows)), 1, tape, C.byref(qs[k]), C.byref(scratch)))
        qs[k].group_size = W
    d_out = torch.zeros((16, W * pitch), dtype=torch.uint8, device="cuda")
    res, outs = {}, {}
    for v1 in (0, 1):
        _capi.check(L.pir_db_set_option(h, b"kernel_v1", v1))
        best = 1e9
        for rep in range(3):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for k in range(16):
                _capi.check(L.pir_answer_batch_device(h, C.byref(qs[k]), 1, C.c_void_p(d_out[k].data_ptr()), sp))
            e1.record(st)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 16)
        res[v1] = best
        outs[v1] = d_out.cpu().numpy().copy()
    same = bool(np.array_equal(outs[0], outs[1]))
    gb = rows * W * pitch / 1e6
    print("2^22 rows x %4d B (slot %d B x group %d): shipped kernel %.4f ms (%.0f GB/s) | round-1 two-launch path %.4f ms (%.0f GB/s) | same bytes: %s" % (
        W * pitch, slot_bytes, W, res[0], gb / res[0], res[1], gb / res[1], same), flush=True)
    L.pir_db_destroy(h)
