"""A small run through every kernel path in one process (compute-sanitizer is closed on the GPU pool, so this is run plain):
single queries back to back on one stream (overlapping launches + look-ahead), host calls, both batch paths, keyword
(fused and walk + scan), expand bits (handle scratch and the static no-alloc scratch), answer-with-bits, Evaluate2P,
a 2-shard server.  Every answer pair must recover its slot.  Usage: python tools/all_paths_probe.py"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pir_b200 import _capi, sharding  # noqa: E402

L = _capi.lib()
rng = np.random.default_rng(11)
LOG = int(os.environ.get("LOG_ROWS", "15"))


def keys(nb, a, kw=False, W=1):
    q0, q1 = _capi.PirQuery(), _capi.PirQuery()
    tape = (C.c_uint8 * 97).from_buffer_copy(rng.integers(0, 256, size=97, dtype=np.uint8).tobytes())
    _capi.check(L.pir_client_gen_two_server(nb, a, 1, tape, C.byref(q0), C.byref(q1)))
    for q in (q0, q1):
        q.is_keyword = 1 if kw else 0
        q.group_size = W
    return q0, q1


def slot(h, idx, S, W=1):
    w = np.zeros(S * W, np.uint8)
    _capi.check(L.pir_db_download(h, idx * W, w.ctypes.data_as(C.c_void_p), W))
    return w


ok = True
for S, W in ((256, 1), (48, 3), (1024, 1)):
    n = (1 << LOG) * W
    h = C.c_void_p()
    _capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
    _capi.check(L.pir_db_fill_synthetic(h, 5))
    nb = L.pir_num_bits_for(n, W, 0)
    rows = n // W
    idxs = [int(x) for x in rng.integers(0, rows, size=6)]
    pairs = [keys(nb, i, W=W) for i in idxs]
    A = S * W
    out = np.zeros((2, len(idxs), A), np.uint8)
    for k, (q0, q1) in enumerate(pairs):          # host calls
        _capi.check(L.pir_answer(h, C.byref(q0), out[0, k].ctypes.data_as(C.c_void_p), A))
        _capi.check(L.pir_answer(h, C.byref(q1), out[1, k].ctypes.data_as(C.c_void_p), A))
        ok &= bool(np.array_equal(out[0, k] ^ out[1, k], slot(h, idxs[k], S, W)))
    for mode in (1, 2):                           # both batch paths
        _capi.check(L.pir_db_set_option(h, b"batch_mode", mode))
        for s in (0, 1):
            arr = (_capi.PirQuery * len(pairs))(*[p[s] for p in pairs])
            _capi.check(L.pir_answer_batch(h, arr, len(pairs), out[s].ctypes.data_as(C.c_void_p), out[s].size))
        ok &= all(np.array_equal(out[0, k] ^ out[1, k], slot(h, idxs[k], S, W)) for k in range(len(idxs)))
    _capi.check(L.pir_db_set_option(h, b"batch_mode", 0))
    bits = np.zeros(rows, np.uint8)               # expand + with-bits
    _capi.check(L.pir_expand_bits(h, C.byref(pairs[0][0]), bits.ctypes.data_as(C.POINTER(C.c_uint8)), rows))
    wb = np.zeros(A, np.uint8)
    _capi.check(L.pir_answer_with_bits(h, W, bits.ctypes.data_as(C.POINTER(C.c_uint8)), rows, wb.ctypes.data_as(C.c_void_p), A))
    ok &= bool(np.array_equal(wb, out[0, 0]))
    if W == 1:                                    # keyword paths
        i = np.arange(n, dtype=np.uint64)
        kws = np.ascontiguousarray((i << np.uint64(32 - LOG)) | ((i * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(64 - (32 - LOG))))
        _capi.check(L.pir_db_set_keywords(h, kws.ctypes.data_as(C.POINTER(C.c_uint64)), n))
        q0, q1 = keys(32, int(kws[idxs[1]]), kw=True)
        for unf in (0, 1):
            _capi.check(L.pir_db_set_option(h, b"force_unfused", unf))
            a0, a1 = np.zeros(A, np.uint8), np.zeros(A, np.uint8)
            _capi.check(L.pir_answer(h, C.byref(q0), a0.ctypes.data_as(C.c_void_p), A))
            _capi.check(L.pir_answer(h, C.byref(q1), a1.ctypes.data_as(C.c_void_p), A))
            ok &= bool(np.array_equal(a0 ^ a1, slot(h, idxs[1], S)))
        _capi.check(L.pir_db_set_option(h, b"force_unfused", 0))
    L.pir_db_destroy(h)
    print("S=%d W=%d ok=%s" % (S, W, ok), flush=True)

# device calls back to back (overlap + look-ahead), the no-alloc expand, Evaluate2P: need torch for device buffers
import torch  # noqa: E402
n, S = 1 << (LOG + 2), 256
h = C.c_void_p()
_capi.check(L.pir_db_create(0, n, S, 0, n, C.byref(h)))
_capi.check(L.pir_db_fill_synthetic(h, 6))
nb = L.pir_num_bits_for(n, 1, 0)
idxs = [int(x) for x in rng.integers(0, n, size=10)]
pairs = [keys(nb, i) for i in idxs]
d_out = torch.zeros((20, S), dtype=torch.uint8, device="cuda")
st = torch.cuda.Stream()
for k, (q0, q1) in enumerate(pairs):
    _capi.check(L.pir_answer_batch_device(h, C.byref(q0), 1, C.c_void_p(d_out[2 * k].data_ptr()), C.c_void_p(st.cuda_stream)))
    _capi.check(L.pir_answer_batch_device(h, C.byref(q1), 1, C.c_void_p(d_out[2 * k + 1].data_ptr()), C.c_void_p(st.cuda_stream)))
torch.cuda.synchronize()
got = d_out.cpu().numpy()
ok &= all(np.array_equal(got[2 * k] ^ got[2 * k + 1], slot(h, idxs[k], S)) for k in range(len(idxs)))
d_bits = torch.empty(n, dtype=torch.uint8, device="cuda")
_capi.check(L.pir_expand_bits_noalloc(h, C.byref(pairs[0][0]), C.c_void_p(d_bits.data_ptr()), n, None))
torch.cuda.synchronize()
bits = np.zeros(n, np.uint8)
_capi.check(L.pir_expand_bits(h, C.byref(pairs[0][0]), bits.ctypes.data_as(C.POINTER(C.c_uint8)), n))
ok &= bool(np.array_equal(bits, d_bits.cpu().numpy()))
xs = np.ascontiguousarray(rng.integers(0, n, size=500, dtype=np.uint64))
xs[0] = idxs[0]
v = [np.zeros(500, np.int64), np.zeros(500, np.int64)]
for s in (0, 1):
    _capi.check(L.pir_dpf_eval_points(0, C.byref(pairs[0][s]), xs.ctypes.data_as(C.POINTER(C.c_uint64)), 500, v[s].ctypes.data_as(C.POINTER(C.c_int64))))
ok &= bool(np.array_equal(v[0] + v[1], (xs == idxs[0]).astype(np.int64)))
L.pir_db_destroy(h)
print("device calls / no-alloc / Evaluate2P ok=%s" % ok, flush=True)
# server of 2 shards on one GPU and a combine ring
with sharding.Server([0, 0], n, S) as srv:
    srv.fill_synthetic(6)
    for k in range(4):
        ok &= bool(np.array_equal(srv.answer(pairs[k][0]) ^ srv.answer(pairs[k][1]), got[2 * k] ^ got[2 * k + 1]))
    b = srv.answer_batch([pairs[0][0], pairs[1][0], pairs[2][0]])
    ok &= bool(np.array_equal(b[1], got[2]))
print("server ok=%s" % ok, flush=True)
print("ALL OK" if ok else "MISMATCH", flush=True)
sys.exit(0 if ok else 1)
