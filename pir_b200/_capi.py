"""ctypes binding of include/pir_b200.h (the C ABI a cgo shim would bind).

The shared library is built in-tree at pir_b200/lib/libpir_b200.so by pir_b200/csrc/Makefile.
There is no fallback: if the library is missing, importing this module raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libpir_b200.so")
DEV_LIB_PATH = os.path.join(_HERE, "lib", "libpir_b200_dev.so")   # `make dev`: instrumentation for tools/ only
CSRC_DIR = os.path.join(_HERE, "csrc")

PIR_MAX_NUM_BITS = 64
PIR_CW_BYTES = 18
PIR_NUM_PRF_KEYS = 4

PIR_OK = 0
PIR_ERR_BAD_ARG = -1
PIR_ERR_BAD_KEY = -2
PIR_ERR_CUDA = -3
PIR_ERR_OOM = -4
PIR_ERR_UNSUPPORTED = -5

# every symbol include/pir_b200.h declares
EXPORTS = [
    "pir_abi_version", "pir_last_error", "pir_device_count",
    "pir_db_create", "pir_db_destroy", "pir_db_upload", "pir_db_download", "pir_db_fill_synthetic",
    "pir_db_set_keywords", "pir_db_set_row_keywords", "pir_db_slot_pitch", "pir_db_device", "pir_db_rows", "pir_db_set_option",
    "pir_answer", "pir_answer_batch", "pir_expand_bits", "pir_expand_bits_noalloc", "pir_answer_with_bits",
    "pir_answer_batch_device", "pir_dpf_eval_points", "pir_xor_combine", "pir_num_bits_for",
    "pir_kernel_launch_count", "pir_client_gen_two_server",
    "pir_query_wire_size", "pir_query_encode", "pir_query_decode",
    "pir_db_set_kernel_timing", "pir_db_get_kernel_timing", "pir_db_time_expansion",
    "pir_combine_wait_read_reset", "pir_combine_wait_device",
    "pir_server_create", "pir_server_destroy", "pir_server_num_shards", "pir_server_shard",
    "pir_server_upload", "pir_server_fill_synthetic", "pir_server_set_keywords", "pir_server_set_option", "pir_server_answer", "pir_server_answer_batch",
    "pir_combine_create", "pir_combine_export", "pir_combine_open", "pir_combine_destroy",
    "pir_combine_reset", "pir_answer_accumulate_device", "pir_combine_wait_read", "pir_combine_slot_ptr",
    "pir_audit_shared_query", "pir_audit_shared_query_with_bits",
]
PIR_IPC_HANDLE_BYTES = 64


class PirQuery(C.Structure):
    """`pir_query` POD: QueryShare (query.go:14-22) + Key2P (dpf/common.go:30-35) + PrfKeys."""

    _fields_ = [
        ("prf_keys", (C.c_uint8 * 16) * PIR_NUM_PRF_KEYS),
        ("s_init", C.c_uint8 * 16),
        ("t_init", C.c_uint8),
        ("share_number", C.c_uint8),
        ("is_keyword", C.c_uint8),
        ("is_two_party", C.c_uint8),
        ("num_bits", C.c_uint32),
        ("final_cw", C.c_int64),
        ("group_size", C.c_uint32),
        ("reserved", C.c_uint32),
        ("cw", (C.c_uint8 * PIR_CW_BYTES) * PIR_MAX_NUM_BITS),
    ]

    def to_bytes(self):
        return bytes(memoryview(self).cast("B"))

    @classmethod
    def from_bytes(cls, b):
        q = cls()
        if len(b) != C.sizeof(cls):
            raise ValueError("pir_query is %d bytes, got %d" % (C.sizeof(cls), len(b)))
        C.memmove(C.byref(q), bytes(b), len(b))
        return q


class PirError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pir_b200 error %d: %s" % (code, msg))
        self.code = code


def use_dev_library():
    """tools/ only: load the -DPIRB_DEV build (in-kernel phase counters, PIRB_* environment switches)."""
    global LIB_PATH, _lib
    if _lib is not None:
        raise RuntimeError("the library is already loaded")
    if not os.path.exists(DEV_LIB_PATH):
        subprocess.check_call(["make", "-C", CSRC_DIR, "-s", "-j8", "dev"])
    LIB_PATH = DEV_LIB_PATH


def build(force=False, extra=""):
    """Compile the CUDA library for sm_100a (nvcc cross-compiles without a GPU)."""
    args = ["make", "-C", CSRC_DIR, "-s", "-j8"]
    if force:
        args.append("-B")
    if extra:
        args.append("EXTRA=" + extra)
    subprocess.check_call(args)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "pir_b200: %s is missing — run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u8p, u64p = C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_uint64)
    qp = C.POINTER(PirQuery)
    L.pir_abi_version.restype = C.c_int
    L.pir_last_error.restype = C.c_char_p
    L.pir_device_count.restype = C.c_int
    L.pir_db_create.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.POINTER(vp)]
    L.pir_db_destroy.argtypes = [vp]
    L.pir_db_upload.argtypes = [vp, C.c_uint64, vp, C.c_uint64]
    L.pir_db_download.argtypes = [vp, C.c_uint64, vp, C.c_uint64]
    L.pir_db_fill_synthetic.argtypes = [vp, C.c_uint64]
    L.pir_db_set_keywords.argtypes = [vp, u64p, C.c_uint64]
    L.pir_db_slot_pitch.argtypes = [vp]
    L.pir_db_slot_pitch.restype = C.c_uint32
    L.pir_db_device.argtypes = [vp]
    L.pir_db_rows.argtypes = [vp, C.c_uint32, u64p, u64p]
    L.pir_db_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.pir_db_set_row_keywords.argtypes = [vp, C.c_uint32, u64p, C.c_uint64]
    L.pir_server_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.pir_answer.argtypes = [vp, qp, vp, C.c_size_t]
    L.pir_answer_batch.argtypes = [vp, qp, C.c_uint32, vp, C.c_size_t]
    L.pir_expand_bits.argtypes = [vp, qp, u8p, C.c_uint64]
    L.pir_expand_bits_noalloc.argtypes = [vp, qp, vp, C.c_uint64, vp]
    L.pir_answer_with_bits.argtypes = [vp, C.c_uint32, u8p, C.c_uint64, vp, C.c_size_t]
    L.pir_answer_batch_device.argtypes = [vp, qp, C.c_uint32, vp, vp]
    L.pir_dpf_eval_points.argtypes = [C.c_int, qp, u64p, C.c_uint64, C.POINTER(C.c_int64)]
    L.pir_xor_combine.argtypes = [vp, vp, C.c_size_t]
    L.pir_xor_combine.restype = None
    L.pir_num_bits_for.argtypes = [C.c_uint64, C.c_uint32, C.c_int]
    L.pir_num_bits_for.restype = C.c_uint32
    L.pir_kernel_launch_count.restype = C.c_uint64
    L.pir_query_wire_size.argtypes = [C.c_uint32]
    L.pir_query_wire_size.restype = C.c_size_t
    L.pir_query_encode.argtypes = [qp, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    L.pir_query_decode.argtypes = [vp, C.c_size_t, qp]
    L.pir_client_gen_two_server.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64, u8p, qp, qp]
    L.pir_db_set_kernel_timing.argtypes = [vp, C.c_int]
    L.pir_db_get_kernel_timing.argtypes = [vp, C.POINTER(C.c_double), u64p]
    L.pir_server_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(vp)]
    L.pir_server_destroy.argtypes = [vp]
    L.pir_server_num_shards.argtypes = [vp]
    L.pir_server_shard.argtypes = [vp, C.c_int, C.POINTER(vp)]
    L.pir_server_upload.argtypes = [vp, C.c_uint64, vp, C.c_uint64]
    L.pir_server_fill_synthetic.argtypes = [vp, C.c_uint64]
    L.pir_server_set_keywords.argtypes = [vp, u64p, C.c_uint64]
    L.pir_server_answer.argtypes = [vp, qp, vp, C.c_size_t]
    L.pir_server_answer_batch.argtypes = [vp, qp, C.c_uint32, vp, C.c_size_t]
    L.pir_combine_create.argtypes = [C.c_int, C.c_uint32, C.c_size_t, C.POINTER(vp)]
    L.pir_combine_export.argtypes = [vp, u8p]
    L.pir_combine_open.argtypes = [C.c_int, u8p, C.c_uint32, C.c_size_t, C.POINTER(vp)]
    L.pir_combine_destroy.argtypes = [vp]
    L.pir_combine_reset.argtypes = [vp, C.c_uint32, C.c_uint32]
    L.pir_answer_accumulate_device.argtypes = [vp, qp, vp, C.c_uint32, vp]
    L.pir_combine_wait_read.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_size_t, C.c_uint32]
    L.pir_combine_wait_read_reset.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_size_t, C.c_uint32]
    L.pir_combine_wait_device.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp]
    L.pir_db_time_expansion.argtypes = [vp, qp, C.c_uint32, C.POINTER(C.c_double)]
    L.pir_combine_slot_ptr.argtypes = [vp, C.c_uint32]
    L.pir_combine_slot_ptr.restype = vp
    L.pir_audit_shared_query.argtypes = [vp, qp, vp, C.c_size_t, vp, C.c_size_t]
    L.pir_audit_shared_query_with_bits.argtypes = [vp, C.c_uint32, u8p, C.c_uint64, vp, C.c_size_t, vp, C.c_size_t]
    L.pir_debug_round_keys.argtypes = [u8p, C.POINTER(C.c_uint32)]
    L.pir_debug_t0.argtypes = [C.POINTER(C.c_uint32)]
    if L.pir_abi_version() != 1:
        raise ImportError("pir_b200: ABI version mismatch")
    _lib = L
    return L


def check(rc):
    if rc != PIR_OK:
        raise PirError(rc, lib().pir_last_error().decode("utf-8", "replace"))
