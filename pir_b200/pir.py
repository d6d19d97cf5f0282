"""Mirror of the reference's `pir` package for the secret-shared (two-server DPF) query path.

Same names, argument meaning and error behaviour as db.go / slot.go / query.go / testing.go for that
path; the server methods run on the GPU through the C ABI (include/pir_b200.h).  `nprocs` is accepted
and ignored, like the reference ignores it in PrivateSecretSharedQueryWithExpandedBits (db.go:75).
"""
import ctypes as C
import math
import os

import numpy as np

from . import _capi, dpf


class Slot:
    """slot.go:14-16"""

    def __init__(self, Data=b""):
        self.Data = bytearray(Data)

    def Equal(self, other):  # slot.go:33-43
        if other is None:
            return False
        return bytes(self.Data) == bytes(other.Data)

    def Compare(self, other):  # slot.go:56-58 (bytes.Compare: -1, 0, 1)
        a, b = bytes(self.Data), bytes(other.Data)
        return (a > b) - (a < b)

    def ToString(self):  # slot.go:61-63: trailing zero bytes removed (an all-zero slot keeps one)
        raw = bytes(self.Data).rstrip(b"\x00")
        return (raw if raw else b"\x00").decode("latin-1")

    def __repr__(self):
        return "Slot(%s)" % bytes(self.Data).hex()


def XorSlots(a, b):
    """slot.go:19-30: a ^= b over min(len a, len b) bytes."""
    n = min(len(a.Data), len(b.Data))
    if n:
        x = np.frombuffer(bytes(a.Data[:n]), dtype=np.uint8) ^ np.frombuffer(bytes(b.Data[:n]), dtype=np.uint8)
        a.Data[:n] = x.tobytes()


def NewSlot(b):  # slot.go:148-152
    return Slot(b)


def NewSlotFromString(s, slotSize):
    """slot.go:137-145: the string's bytes, zero padded up to slotSize (never truncated)."""
    raw = s.encode("latin-1") if isinstance(s, str) else bytes(s)
    return Slot(raw + bytes(max(0, slotSize - len(raw))))


def GetRequiredSlotSize(data):
    """slot.go:174-186: the longest string, in bytes."""
    return max([len(s.encode("latin-1") if isinstance(s, str) else bytes(s)) for s in data], default=0)


def NewEmptySlot(numBytes):
    return Slot(bytes(numBytes))


def NewRandomSlot(numBytes):
    return Slot(os.urandom(numBytes))


class SecretSharedQueryResult:
    """db.go:29-32"""

    def __init__(self, SlotBytes, Shares):
        self.SlotBytes = SlotBytes
        self.Shares = Shares


class QueryShare:
    """query.go:14-22"""

    def __init__(self):
        self.KeyTwoParty = None
        self.KeyMultiParty = None
        self.PrfKeys = []
        self.IsKeywordBased = False
        self.IsTwoParty = True
        self.ShareNumber = 0
        self.GroupSize = 1

    def to_pod(self, num_bits=None):
        if not self.IsTwoParty:
            raise NotImplementedError("multi-party shares are not implemented (dpf/client.go:152-156)")
        return dpf.pod_from_parts(self.PrfKeys, self.KeyTwoParty, self.ShareNumber, num_bits,
                                  group_size=self.GroupSize, is_keyword=self.IsKeywordBased)

    # The reference keeps shares in process and has no serialisation; these two carry Go's encoding.BinaryMarshaler /
    # BinaryUnmarshaler names and use the wire layout of include/pir_b200.h (pir_query_encode / pir_query_decode).
    def MarshalBinary(self):
        L = _capi.lib()
        q = self.to_pod()
        buf = (C.c_uint8 * L.pir_query_wire_size(q.num_bits))()
        n = C.c_size_t()
        _capi.check(L.pir_query_encode(C.byref(q), buf, len(buf), C.byref(n)))
        return bytes(buf[:n.value])

    @classmethod
    def UnmarshalBinary(cls, data):
        L = _capi.lib()
        q = _capi.PirQuery()
        raw = (C.c_uint8 * max(len(data), 1)).from_buffer_copy(bytes(data) or b"\0")
        _capi.check(L.pir_query_decode(raw, len(data), C.byref(q)))
        s = cls()
        s.PrfKeys = [dpf.PrfKey(bytes(q.prf_keys[i])) for i in range(dpf.initPRFLen)]
        s.KeyTwoParty = dpf.Key2P(bytes(q.s_init), q.t_init, [bytes(q.cw[i]) for i in range(q.num_bits)], q.final_cw)
        s.IsKeywordBased = bool(q.is_keyword)
        s.IsTwoParty = bool(q.is_two_party)
        s.ShareNumber = q.share_number
        s.GroupSize = q.group_size
        return s


class DBMetadata:
    """db.go:15-18 and the client-side query constructors of query.go:42-101."""

    def __init__(self, SlotBytes=0, DBSize=0):
        self.SlotBytes = SlotBytes
        self.DBSize = DBSize

    def NewIndexQueryShares(self, index, groupSize, numShares, _tape=None):
        return self.newQueryShares(index, groupSize, numShares, True, _tape=_tape)

    def NewKeywordQueryShares(self, keyword, groupSize, numShares, _tape=None):
        return self.newQueryShares(keyword, groupSize, numShares, False, _tape=_tape)

    def newQueryShares(self, key, groupSize, numShares, isIndexQuery, _tape=None):
        dimHeight = self.DBSize // groupSize
        if dimHeight == 0:
            raise RuntimeError("database height is set to zero; something is wrong")  # query.go:56-58
        numBits = dimHeight.bit_length()  # uint(math.Log2(float64(dimHeight)) + 1)
        if not isIndexQuery:
            numBits = 32
        pf = dpf.ClientInitialize(numBits, _tape=_tape)
        if numShares != 2:
            pf.GenerateMultiServer(key, 1, numShares)
        keys = pf.GenerateTwoServer(key, 1, _tape=_tape)
        # the reference checks the range only for... every query kind, after key generation (query.go:79-81)
        if key >= dimHeight:
            raise RuntimeError("requesting key outside of domain")
        shares = []
        for i in range(numShares):
            s = QueryShare()
            s.ShareNumber = i
            s.PrfKeys = pf.PrfKeys
            s.IsKeywordBased = not isIndexQuery
            s.GroupSize = groupSize
            s.KeyTwoParty = keys[i]
            s.IsTwoParty = True
            shares.append(s)
        return shares

    def IndexToCoordinates(self, index, width, height):  # db.go:399-401
        return index // width, index % width

    def GetDimentionsForDatabase(self, height, groupSize):
        """db.go:406-420 (integer divisions inside math.Ceil, as written there)."""
        dimWidth = self.DBSize // (height * groupSize)
        if dimWidth == 0:
            dimWidth = 1
        dimHeight = self.DBSize // (dimWidth * groupSize)
        return dimWidth * groupSize, dimHeight

    def GetSqrtOfDBSize(self):  # db.go:423-425
        return int(math.sqrt(float(self.DBSize)) + 1)


class Database(DBMetadata):
    """db.go:22-26.  Slots live on the host like in the reference; Upload() (or the first query) makes
    the flat device copy behind a pir_db handle.  Mutating Slots afterwards needs another Upload()."""

    def __init__(self, device=0):
        super().__init__()
        self.Slots = []
        self.Keywords = None
        self.device = device
        self._h = None
        self._shard = None

    # -- handle management (no counterpart in the reference: its DB is host memory) --
    def Upload(self, slot_begin=0, slot_count=None):
        self.Close()
        L = _capi.lib()
        if slot_count is None:
            slot_count = self.DBSize - slot_begin
        h = C.c_void_p()
        _capi.check(L.pir_db_create(self.device, self.DBSize, self.SlotBytes, slot_begin, slot_count, C.byref(h)))
        self._h = h
        self._shard = (slot_begin, slot_count)
        if self.Slots:
            flat = np.zeros((slot_count, self.SlotBytes), dtype=np.uint8)
            for i in range(slot_count):
                d = self.Slots[slot_begin + i].Data
                flat[i, :len(d)] = np.frombuffer(bytes(d[:self.SlotBytes]), dtype=np.uint8)
            _capi.check(L.pir_db_upload(h, slot_begin, flat.ctypes.data_as(C.c_void_p), slot_count))
        if self.Keywords is not None:
            kw = np.ascontiguousarray(self.Keywords[slot_begin:slot_begin + slot_count], dtype=np.uint64)
            _capi.check(L.pir_db_set_keywords(h, kw.ctypes.data_as(C.POINTER(C.c_uint64)), kw.size))
        return self

    def UploadArray(self, flat, slot_begin=0):
        """Upload a dense (n, SlotBytes) uint8 array without building Slot objects."""
        flat = np.ascontiguousarray(flat, dtype=np.uint8).reshape(-1, self.SlotBytes)
        if self._h is None:
            self.Upload(slot_begin, flat.shape[0]) if not self.Slots else self.Upload()
        _capi.check(_capi.lib().pir_db_upload(self._h, slot_begin, flat.ctypes.data_as(C.c_void_p), flat.shape[0]))
        return self

    def FillSynthetic(self, seed):
        if self._h is None:
            self.Upload()
        _capi.check(_capi.lib().pir_db_fill_synthetic(self._h, seed))
        return self

    def Close(self):
        if self._h is not None:
            _capi.lib().pir_db_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass

    def _handle(self):
        if self._h is None:
            self.Upload()
        return self._h

    def _rows(self, groupSize):
        a, b = C.c_uint64(), C.c_uint64()
        _capi.check(_capi.lib().pir_db_rows(self._handle(), groupSize, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- db.go:68-72 --
    def PrivateSecretSharedQuery(self, query, nprocs=1):
        q = query.to_pod(num_bits=len(query.KeyTwoParty.CW))
        n = query.GroupSize * self.SlotBytes
        out = (C.c_uint8 * n)()
        _capi.check(_capi.lib().pir_answer(self._handle(), C.byref(q), out, n))
        return self._result(bytes(out), query.GroupSize)

    def PrivateSecretSharedQueryBatch(self, queries, nprocs=1):
        """Many queries in one call (no counterpart in the reference)."""
        arr = (_capi.PirQuery * len(queries))(*[q.to_pod(num_bits=len(q.KeyTwoParty.CW)) for q in queries])
        n = queries[0].GroupSize * self.SlotBytes
        out = (C.c_uint8 * (n * len(queries)))()
        _capi.check(_capi.lib().pir_answer_batch(self._handle(), arr, len(queries), out, n * len(queries)))
        raw = bytes(out)
        return [self._result(raw[i * n:(i + 1) * n], queries[0].GroupSize) for i in range(len(queries))]

    # -- db.go:110-174 --
    def ExpandSharedQuery(self, query, nprocs=1):
        q = query.to_pod(num_bits=len(query.KeyTwoParty.CW))
        _, n_rows = self._rows(query.GroupSize)
        bits = np.zeros(max(n_rows, 1), dtype=np.uint8)
        _capi.check(_capi.lib().pir_expand_bits(self._handle(), C.byref(q), bits.ctypes.data_as(C.POINTER(C.c_uint8)), n_rows))
        return [bool(b) for b in bits[:n_rows]]

    # -- db.go:75-107 --
    def PrivateSecretSharedQueryWithExpandedBits(self, query, bits, nprocs=1):
        b = np.ascontiguousarray(np.asarray(bits, dtype=np.uint8))
        n = query.GroupSize * self.SlotBytes
        out = (C.c_uint8 * n)()
        _capi.check(_capi.lib().pir_answer_with_bits(self._handle(), query.GroupSize,
                                                     b.ctypes.data_as(C.POINTER(C.c_uint8)), b.size, out, n))
        return self._result(bytes(out), query.GroupSize)

    def _result(self, raw, groupSize):
        S = self.SlotBytes
        return SecretSharedQueryResult(S, [Slot(raw[c * S:(c + 1) * S]) for c in range(groupSize)])

    # -- db.go:364-394 --
    def BuildForData(self, data):
        self.BuildForDataWithSlotSize(data, GetRequiredSlotSize(data))

    def BuildForDataWithSlotSize(self, data, slotSize):
        self.Close()
        self.SlotBytes = slotSize
        self.DBSize = len(data)
        self.Slots = []
        for s in data:
            d = bytearray(slotSize)
            raw = s.encode("latin-1") if isinstance(s, str) else bytes(s)
            d[:min(len(raw), slotSize)] = raw[:slotSize]
            self.Slots.append(Slot(d))

    def SetKeywords(self, keywords):
        self.Keywords = list(keywords)
        if self._h is not None:
            self.Upload(*self._shard)


def NewDatabase():
    return Database()


def GetOptimalDBDimentions(slotSize, dbSize):
    """db.go:428-434"""
    height = int(max(1.0, math.sqrt(float(dbSize * slotSize))))
    return int(math.ceil(float(dbSize) / float(height))), height


def GetOptimalWeightedDBDimentions(slotSize, dbSize, weight):
    """db.go:438-446"""
    width, height = GetOptimalDBDimentions(slotSize, dbSize)
    return width // weight, height * weight


def GenerateRandomDB(size, numBytes, device=0):
    """testing.go:5-17"""
    db = Database(device)
    db.SlotBytes = numBytes
    db.DBSize = size
    db.Slots = [NewRandomSlot(numBytes) for _ in range(size)]
    return db


def GenerateEmptyDB(size, numBytes, device=0):
    """testing.go:20-36"""
    db = Database(device)
    db.SlotBytes = numBytes
    db.DBSize = size
    db.Slots = [NewEmptySlot(numBytes) for _ in range(size)]
    return db


def Recover(resShares):
    """query.go:278-297: XOR of the servers' result shares."""
    numSlots = len(resShares[0].Shares)
    res = [NewEmptySlot(resShares[0].SlotBytes) for _ in range(numSlots)]
    for share in resShares:
        for j in range(numSlots):
            XorSlots(res[j], share.Shares[j])
    return res
