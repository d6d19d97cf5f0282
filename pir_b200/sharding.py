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
