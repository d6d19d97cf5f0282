"""Row sharding of one PIR server's database over the GPUs of a box, and the XOR combine.

SURVEY.md 8(e): the answer is an XOR over rows (db.go:91-104), so GPU g of G holds a contiguous range
of rows and produces a partial answer of GroupSize*SlotBytes bytes; the server's answer is the XOR
of the G partials.  No database byte ever crosses GPUs and no NCCL collective is involved (NCCL has no
XOR reduction): partials are gathered (torch.distributed object/tensor gather on whatever backend the
process group has, or peer copies) and folded with pir_xor_combine (slot.go:19-30 semantics).
"""
import ctypes as C

import numpy as np

from . import _capi


def shard_slots(db_size, group_size, rank, world):
    """Slot range [begin, begin+count) of `rank`: whole rows (group_size slots each), contiguous, the
    ragged tail (db_size % group_size slots, never read by db.go:79,114) goes to the last rank."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d outside world %d" % (rank, world))
    if group_size < 1:
        raise ValueError("group_size must be >= 1")
    rows = db_size // group_size
    r0 = rank * rows // world
    r1 = (rank + 1) * rows // world
    begin = r0 * group_size
    end = db_size if rank == world - 1 else r1 * group_size
    return begin, end - begin


def shard_rows(db_size, group_size, rank, world):
    """Row range [first, first+count) matching shard_slots (what pir_db_rows reports for the shard)."""
    rows = db_size // group_size
    r0 = rank * rows // world
    r1 = (rank + 1) * rows // world
    return r0, r1 - r0


def xor_combine(partials):
    """XOR-fold equal-length byte strings / uint8 arrays with the library's pir_xor_combine."""
    L = _capi.lib()
    it = iter(partials)
    acc = np.array(np.frombuffer(bytes(next(it)), dtype=np.uint8))
    for p in it:
        b = np.ascontiguousarray(np.frombuffer(bytes(p), dtype=np.uint8))
        if b.size != acc.size:
            raise ValueError("partial answers differ in length: %d vs %d" % (b.size, acc.size))
        L.pir_xor_combine(acc.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), acc.size)
    return acc


def gather_and_combine(partial, dist, dst=0):
    """Host reduction over a torch.distributed process group: every rank passes its partial answer;
    rank `dst` gets the combined answer (others get None)."""
    world = dist.get_world_size()
    parts = [None] * world if dist.get_rank() == dst else None
    dist.gather_object(bytes(partial), parts, dst=dst)
    if parts is None:
        return None
    return xor_combine(parts)


class Server:
    """One PIR server whose database is row-sharded over several GPUs, driven from ONE process
    (pir_server_* in include/pir_b200.h).  Every shard's kernel XORs its partial answer into a single
    accumulator on devices[0] (peer stores over NVLink); there is no combine pass and no NCCL."""

    def __init__(self, devices, db_size, slot_bytes, shard_align=1):
        self.L = _capi.lib()
        self.db_size, self.slot_bytes = db_size, slot_bytes
        arr = (C.c_int * len(devices))(*devices)
        self.h = C.c_void_p()
        _capi.check(self.L.pir_server_create(arr, len(devices), db_size, slot_bytes, shard_align, C.byref(self.h)))

    @property
    def num_shards(self):
        return self.L.pir_server_num_shards(self.h)

    def shard(self, i):
        db = C.c_void_p()
        _capi.check(self.L.pir_server_shard(self.h, i, C.byref(db)))
        return db

    def upload(self, dense, slot_index=0):
        dense = np.ascontiguousarray(dense, dtype=np.uint8)
        _capi.check(self.L.pir_server_upload(self.h, slot_index, dense.ctypes.data_as(C.c_void_p),
                                             dense.size // self.slot_bytes))

    def fill_synthetic(self, seed):
        _capi.check(self.L.pir_server_fill_synthetic(self.h, seed))

    def set_keywords(self, keywords):
        """Database.SetKeywords (db.go:392-394): one keyword per slot; every shard keeps its own slots' keywords."""
        kw = np.ascontiguousarray(keywords, dtype=np.uint64)
        _capi.check(self.L.pir_server_set_keywords(self.h, kw.ctypes.data_as(C.POINTER(C.c_uint64)), kw.size))

    def set_option(self, name, value):
        """pir_server_set_option: path selection on every shard (results are bit-identical for every setting)."""
        _capi.check(self.L.pir_server_set_option(self.h, name.encode(), int(value)))

    def answer(self, q):
        n = q.group_size * self.slot_bytes
        out = np.zeros(n, dtype=np.uint8)
        _capi.check(self.L.pir_server_answer(self.h, C.byref(q), out.ctypes.data_as(C.c_void_p), n))
        return out

    def answer_batch(self, qs):
        n = qs[0].group_size * self.slot_bytes
        arr = (_capi.PirQuery * len(qs))(*qs)
        out = np.zeros(n * len(qs), dtype=np.uint8)
        _capi.check(self.L.pir_server_answer_batch(self.h, arr, len(qs), out.ctypes.data_as(C.c_void_p), out.size))
        return out.reshape(len(qs), n)

    def close(self):
        if self.h:
            self.L.pir_server_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Combine:
    """Ring of combine slots for the one-process-per-GPU launch (pir_combine_* in include/pir_b200.h).
    The owner rank creates and exports; the others open the CUDA IPC handle."""

    def __init__(self, device, n_slots, slot_stride, handle=None):
        self.L = _capi.lib()
        self.n_slots, self.stride, self.owner = n_slots, slot_stride, handle is None
        self.h = C.c_void_p()
        if handle is None:
            _capi.check(self.L.pir_combine_create(device, n_slots, slot_stride, C.byref(self.h)))
        else:
            buf = (C.c_uint8 * _capi.PIR_IPC_HANDLE_BYTES).from_buffer_copy(bytes(handle))
            _capi.check(self.L.pir_combine_open(device, buf, n_slots, slot_stride, C.byref(self.h)))

    def export(self):
        buf = (C.c_uint8 * _capi.PIR_IPC_HANDLE_BYTES)()
        _capi.check(self.L.pir_combine_export(self.h, buf))
        return bytes(buf)

    def reset(self, first=0, count=None):
        _capi.check(self.L.pir_combine_reset(self.h, first, self.n_slots - first if count is None else count))

    def accumulate(self, db, q, slot, stream=None):
        _capi.check(self.L.pir_answer_accumulate_device(db, C.byref(q), self.h, slot, stream))

    def wait_read(self, slot, expected, nbytes, timeout_ms=10000):
        out = np.zeros(nbytes, dtype=np.uint8)
        _capi.check(self.L.pir_combine_wait_read(self.h, slot, expected, out.ctypes.data_as(C.c_void_p), nbytes,
                                                 timeout_ms))
        return out

    def wait_read_reset(self, slot, expected, nbytes, timeout_ms=10000):
        """wait_read, then the slot is zeroed behind the copy (one synchronisation): free for its next query"""
        out = np.zeros(nbytes, dtype=np.uint8)
        _capi.check(self.L.pir_combine_wait_read_reset(self.h, slot, expected, out.ctypes.data_as(C.c_void_p), nbytes,
                                                       timeout_ms))
        return out

    def wait_device(self, slot, expected, stream=None, timeout_ms=10000):
        _capi.check(self.L.pir_combine_wait_device(self.h, slot, expected, timeout_ms, stream))

    def close(self):
        if self.h:
            self.L.pir_combine_destroy(self.h)
            self.h = None


def open_combine_over_group(dist, device, n_slots, slot_stride, owner=0):
    """Rendezvous of the IPC handle over an existing torch.distributed group (plumbing only: 64 bytes,
    once).  Returns this rank's Combine."""
    rank = dist.get_rank()
    comb = Combine(device, n_slots, slot_stride) if rank == owner else None
    box = [comb.export() if rank == owner else None]
    dist.broadcast_object_list(box, src=owner)
    if rank != owner:
        comb = Combine(device, n_slots, slot_stride, handle=box[0])
    return comb
