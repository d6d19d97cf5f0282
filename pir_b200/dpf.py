"""Mirror of the reference's `dpf` package surface for the two-party path, over the C ABI.

Names and argument meaning follow dpf/common.go, dpf/client.go and dpf/server.go; evaluation runs on
the GPU (pir_dpf_eval_points), key generation on the host (pir_client_gen_two_server).
"""
import ctypes as C

import numpy as np

from . import _capi

initPRFLen = 4  # dpf/common.go:10


class PrfKey:
    """dpf/common.go:13-15"""

    def __init__(self, Bytes=b""):
        self.Bytes = bytes(Bytes)


class Key2P:
    """dpf/common.go:30-35"""

    def __init__(self, SInit=b"", TInit=0, CW=None, FinalCW=0):
        self.SInit = bytes(SInit)
        self.TInit = int(TInit)
        self.CW = [bytes(c) for c in (CW or [])]
        self.FinalCW = int(FinalCW)


class Dpf:
    """dpf/common.go:18-27 (only the fields the two-party path uses)."""

    def __init__(self, NumBits, PrfKeys=None, device=0):
        self.NumBits = int(NumBits)
        self.PrfKeys = PrfKeys or []
        self.N = 64
        self.M = 4
        self.device = device

    # ---- client: dpf/client.go:56-150 ----
    def GenerateTwoServer(self, a, b, _tape=None):
        q0, q1 = _capi.PirQuery(), _capi.PirQuery()
        tape = None
        if _tape is not None:
            tape = (C.c_uint8 * 97).from_buffer_copy(bytes(_tape))
        elif self.PrfKeys:
            # keep the PRF keys ClientInitialize drew; fresh randomness for the seeds
            import os
            tape = (C.c_uint8 * 97).from_buffer_copy(b"".join(k.Bytes for k in self.PrfKeys) + os.urandom(33))
        _capi.check(_capi.lib().pir_client_gen_two_server(self.NumBits, int(a), int(b), tape, C.byref(q0), C.byref(q1)))
        if not self.PrfKeys:
            self.PrfKeys = [PrfKey(bytes(q0.prf_keys[i])) for i in range(initPRFLen)]
        return [_key_from_pod(q0), _key_from_pod(q1)]

    def GenerateMultiServer(self, a, b, num_p):
        raise NotImplementedError("not implemented")  # dpf/client.go:152-156 panics the same way

    # ---- server: dpf/server.go:55-101 ----
    def Evaluate2P(self, serverNum, k, x):
        return int(self.Evaluate2PBatch(serverNum, k, [x])[0])

    def Evaluate2PBatch(self, serverNum, k, xs):
        """Many points in one launch; out[i] = Evaluate2P(serverNum, k, xs[i]) as wrapping int64."""
        q = pod_from_parts(self.PrfKeys, k, serverNum, self.NumBits, group_size=1, is_keyword=False)
        xs = np.ascontiguousarray(xs, dtype=np.uint64)
        out = np.empty(xs.size, dtype=np.int64)
        _capi.check(_capi.lib().pir_dpf_eval_points(
            self.device, C.byref(q), xs.ctypes.data_as(C.POINTER(C.c_uint64)), xs.size,
            out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out


def ClientInitialize(numBits, _tape=None):
    """dpf/client.go:17-48: draws the four AES keys."""
    import os
    keys = bytes(_tape[:64]) if _tape is not None else os.urandom(64)
    return Dpf(numBits, [PrfKey(keys[16 * i:16 * i + 16]) for i in range(initPRFLen)])


def ServerInitialize(prfKeys, numBits, device=0):
    """dpf/server.go:16-46.  The AES key schedules are expanded inside every evaluation call."""
    for k in prfKeys:
        if len(k.Bytes) != 16:
            raise ValueError("crypto/aes: invalid key size %d" % len(k.Bytes))  # aes.NewCipher's error
    return Dpf(numBits, [PrfKey(k.Bytes) for k in prfKeys], device=device)


def _key_from_pod(q):
    return Key2P(bytes(q.s_init), q.t_init, [bytes(q.cw[i]) for i in range(q.num_bits)], q.final_cw)


def pod_from_parts(prf_keys, key2p, share_number, num_bits, group_size=1, is_keyword=False):
    """Flatten (PrfKeys, Key2P, ShareNumber, ...) into the pir_query POD of include/pir_b200.h."""
    q = _capi.PirQuery()
    if len(prf_keys) < 3:
        raise ValueError("need at least 3 PRF keys")
    for i, k in enumerate(prf_keys[:initPRFLen]):
        if len(k.Bytes) != 16:
            raise ValueError("crypto/aes: invalid key size %d" % len(k.Bytes))
        C.memmove(q.prf_keys[i], k.Bytes, 16)
    if len(key2p.SInit) != 16:
        raise ValueError("SInit must be 16 bytes")
    C.memmove(q.s_init, key2p.SInit, 16)
    q.t_init = key2p.TInit
    q.share_number = int(share_number)
    q.is_keyword = 1 if is_keyword else 0
    q.is_two_party = 1
    if len(key2p.CW) > _capi.PIR_MAX_NUM_BITS:
        raise ValueError("too many correction words")
    q.num_bits = len(key2p.CW) if num_bits is None else int(num_bits)
    for i, cw in enumerate(key2p.CW):
        if len(cw) != _capi.PIR_CW_BYTES:
            raise ValueError("correction word %d must be 18 bytes" % i)
        C.memmove(q.cw[i], cw, _capi.PIR_CW_BYTES)
    fcw = int(key2p.FinalCW)
    fcw &= (1 << 64) - 1
    q.final_cw = fcw - (1 << 64) if fcw >> 63 else fcw
    q.group_size = int(group_size)
    return q
