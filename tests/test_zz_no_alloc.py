"""pir_expand_bits_noalloc: the allocation-free variant of pir_expand_bits (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its scratch is a file-scope static device array
shared by every call, so the checks are: the output equals the existing entry point's bit for bit (and the oracle's) on
two calls in a row with DIFFERENT keys -- the second call must not see the first call's bits -- and sizes beyond
PIR_NOALLOC_MAX_ROWS are rejected.  Collected last (test_zz_*)."""
import ctypes as C

import numpy as np
import pytest

import oracle
from pir_b200 import _capi
from tests.helpers import Handle, make_queries

pytestmark = pytest.mark.gpu
PIR_NOALLOC_MAX_ROWS = 1 << 24


def _noalloc_bits(L, h, q, d_out, n_rows, stream):
    d_out.fill_(0xFF)   # poison: every byte must be written by the call
    _capi.check(L.pir_expand_bits_noalloc(h.h, C.byref(q), C.c_void_p(d_out.data_ptr()), n_rows, stream))


@pytest.mark.parametrize("db_size,slot_bytes,W,first,count", [
    (1 << 16, 32, 1, 0, None),       # BASELINE configs[0] shape
    (5000, 48, 1, 0, None),          # non-power-of-two height
    (12345, 64, 2, 0, None),         # group size 2, ragged tail
    (1 << 14, 256, 1, 5000, 7000),   # a shard whose first row is not a multiple of 32
])
def test_noalloc_expand_bits_matches_existing_entry_point_twice_in_a_row(gpu_lib, db_size, slot_bytes, W, first, count):
    import torch
    L = gpu_lib
    rows_all = db_size // W
    with Handle(db_size, slot_bytes, slot_begin=first * W, slot_count=None if count is None else count * W) as h:
        h.fill(7)
        r0, n_rows = h.rows(W)
        d_out = torch.empty(n_rows, dtype=torch.uint8, device="cuda")   # the CALLER's buffer, allocated before the calls
        st = torch.cuda.Stream()
        # two calls in a row, different keys and servers; then the first key again
        keys = []
        for k, idx in enumerate((rows_all // 3, rows_all - 1)):
            os_, qs = make_queries(800 + k, db_size, W, idx)
            keys.append((os_[k & 1], qs[k & 1]))
        keys.append(keys[0])
        prev = None
        for o, q in keys:
            _noalloc_bits(L, h, q, d_out, n_rows, C.c_void_p(st.cuda_stream))
            st.synchronize()
            got = d_out.cpu().numpy()
            want = h.expand_bits(q)                         # the existing entry point
            assert np.array_equal(got, want)
            assert np.array_equal(got, oracle.expand_shared_query(o, db_size, row_begin=r0, n_rows=n_rows))
            if prev is not None:
                assert not np.array_equal(got, prev)        # (the two keys do select different rows)
            prev = got
        # back to back on one stream without a synchronisation in between: stream order keeps the calls apart
        d_a = torch.empty(n_rows, dtype=torch.uint8, device="cuda")
        d_b = torch.empty(n_rows, dtype=torch.uint8, device="cuda")
        _noalloc_bits(L, h, keys[0][1], d_a, n_rows, C.c_void_p(st.cuda_stream))
        _noalloc_bits(L, h, keys[1][1], d_b, n_rows, C.c_void_p(st.cuda_stream))
        st.synchronize()
        assert np.array_equal(d_a.cpu().numpy(), h.expand_bits(keys[0][1]))
        assert np.array_equal(d_b.cpu().numpy(), h.expand_bits(keys[1][1]))
        # NULL stream = the handle's own
        _noalloc_bits(L, h, keys[1][1], d_a, n_rows, None)
        torch.cuda.synchronize()
        assert np.array_equal(d_a.cpu().numpy(), h.expand_bits(keys[1][1]))


def test_noalloc_expand_bits_at_the_maximum_and_beyond(gpu_lib, capi):
    import torch
    L = gpu_lib
    # exactly PIR_NOALLOC_MAX_ROWS rows: accepted, bits equal the existing entry point's and the oracle's (tree expansion)
    n, S = PIR_NOALLOC_MAX_ROWS, 16
    with Handle(n, S) as h:
        d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
        for k in range(2):
            os_, qs = make_queries(900 + k, n, 1, (k + 1) * 5_000_011)
            _noalloc_bits(L, h, qs[k], d_out, n, None)
            torch.cuda.synchronize()
            got = d_out.cpu().numpy()
            assert np.array_equal(got, h.expand_bits(qs[k]))
            assert np.array_equal(got, oracle.expand_tree(os_[k], n, nthreads=8))
    # one row block more: rejected with a clear error, nothing written
    n2 = PIR_NOALLOC_MAX_ROWS + 32
    with Handle(n2, S) as h:
        (_, _), (q0, _) = make_queries(902, n2, 1, 77)
        d_out = torch.zeros(n2, dtype=torch.uint8, device="cuda")
        with pytest.raises(capi.PirError) as e:
            _capi.check(L.pir_expand_bits_noalloc(h.h, C.byref(q0), C.c_void_p(d_out.data_ptr()), n2, None))
        assert e.value.code == capi.PIR_ERR_BAD_ARG and "PIR_NOALLOC_MAX_ROWS" in str(e.value)
        assert h.expand_bits(q0).size == n2              # the allocating entry point still serves that size
    # keyword queries and a wrong row count are rejected too
    with Handle(1024, 16) as h:
        (_, _), (q0, _) = make_queries(903, 1024, 1, 5)
        d_out = torch.zeros(1024, dtype=torch.uint8, device="cuda")
        with pytest.raises(capi.PirError):
            _capi.check(L.pir_expand_bits_noalloc(h.h, C.byref(q0), C.c_void_p(d_out.data_ptr()), 1000, None))
        q0.is_keyword = 1
        with pytest.raises(capi.PirError) as e:
            _capi.check(L.pir_expand_bits_noalloc(h.h, C.byref(q0), C.c_void_p(d_out.data_ptr()), 1024, None))
        assert e.value.code == capi.PIR_ERR_UNSUPPORTED
