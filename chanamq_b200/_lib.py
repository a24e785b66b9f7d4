"""ctypes binding of libchanagpu.so — the same symbols a JNA interface on the Scala side binds
(INTEGRATION.md).  Fails loudly if the library has not been built."""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libchanagpu.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "chanamq_gpu.h")

CHQ_DIRECT, CHQ_FANOUT, CHQ_TOPIC, CHQ_HEADERS = 0, 1, 2, 3
CHQ_MODE_COMPAT, CHQ_MODE_SPEC = 0, 1
CHQ_OK, CHQ_E_ZERO_WORDS, CHQ_E_KEY_TOO_LONG, CHQ_E_OVERFLOW, CHQ_E_CUDA, CHQ_E_ARG, CHQ_E_NOT_COMMITTED, CHQ_E_SCRATCH = (
    0, -1, -2, -3, -4, -5, -6, -7)
CHQ_SLOTS = 4
CHQ_IDS_U32, CHQ_IDS_U24 = 0, 1


class ChqError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("chanamq_gpu status %d: %s" % (code, text))
        self.code = code


class ChqResult(C.Structure):
    _fields_ = [("n", C.c_uint32), ("offsets", C.c_void_p), ("queue_ids", C.c_void_p), ("any_durable", C.c_void_p),
                ("err", C.c_void_p), ("ids_format", C.c_int32)]


_lib = None


def declared_symbols():
    """every function include/chanamq_gpu.h declares"""
    text = open(HEADER_PATH).read()
    return sorted(set(re.findall(r"\b(chq_[a-z_]+)\s*\(", text)))


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ChqError(CHQ_E_CUDA, "libchanagpu.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                                   "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32, u8 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32, C.c_uint8
    sig = {
        "chq_last_error": (C.c_char_p, []),
        "chq_table_create": (i32, [C.POINTER(vp), i32, i32, i32]),
        "chq_table_destroy": (i32, [vp]),
        "chq_bind": (i32, [vp, C.c_char_p, u32, C.c_char_p, u32, u8]),
        "chq_bind_args": (i32, [vp, C.c_char_p, u32, C.c_char_p, u32, u8, C.c_char_p, u32]),
        "chq_bind_many": (i32, [vp, vp, vp, vp, vp, vp, u32]),
        "chq_unbind": (i32, [vp, C.c_char_p, u32, C.c_char_p, u32, u8]),
        "chq_unbind_queue": (i32, [vp, C.c_char_p, u32]),
        "chq_commit": (i32, [vp]),
        "chq_commit_full": (i32, [vp]),
        "chq_n_queues": (i32, [vp, C.POINTER(u32)]),
        "chq_queue_name": (i32, [vp, u32, C.POINTER(vp), C.POINTER(u32)]),
        "chq_route": (i32, [vp, vp, vp, u32, vp, vp, C.POINTER(ChqResult)]),
        "chq_route_submit": (i32, [vp, i32, vp, vp, u32]),
        "chq_route_wait": (i32, [vp, i32, C.POINTER(ChqResult)]),
        "chq_set_ids_format": (i32, [vp, i32]),
        "chq_unpack_u24": (None, [vp, u64, vp]),
        "chq_route_device": (i32, [vp, vp, vp, u32, vp, vp, u64, vp, vp, vp, vp, vp]),
        "chq_route_device_static": (i32, [vp, vp, vp, u32, vp, vp, u64, vp, vp, vp, vp, vp]),
        "chq_declare_queues": (i32, [vp, vp, vp, u32]),
        "chq_merge_slices_device": (i32, [i32, vp, vp, u64, vp, u32, u32, vp, vp, u64, vp, vp, vp, vp]),
        "chq_merge_slices_device2": (i32, [i32, vp, vp, u64, vp, u32, u32, vp, vp, u64, vp, vp, vp, u32, vp]),
        "chq_shard_unique_id": (i32, [vp]),
        "chq_shard_create": (i32, [C.POINTER(vp), vp, i32, i32, i32, i32, i32]),
        "chq_shard_destroy": (i32, [vp]),
        "chq_shard_table": (vp, [vp]),
        "chq_shard_bind_many": (i32, [vp, vp, vp, vp, vp, vp, u32, u32]),
        "chq_shard_commit": (i32, [vp]),
        "chq_shard_is_disjoint": (i32, [vp]),
        "chq_shard_reserve": (i32, [vp, u32, u64, u64]),
        "chq_shard_route_device": (i32, [vp, vp, vp, u32, vp, vp, u64, vp, vp, vp, vp, C.POINTER(u32), C.POINTER(u32), vp]),
        "chq_shard_stage_ms": (i32, [vp, i32, C.POINTER(C.c_double), C.POINTER(u32)]),
        "chq_shard_last_error": (C.c_char_p, []),
        "chq_decode": (i32, [vp, u64, vp, u32, vp, u64, C.POINTER(u64), vp, vp, vp]),
        "chq_decode_device": (i32, [vp, u64, vp, u32, vp, u64, vp, vp, vp, vp, vp, vp]),
        "chq_publish_keys_device": (i32, [vp, vp, u64, vp, vp, vp, vp, vp]),
        "chq_route_keys_device": (i32, [vp, vp, vp, vp, u32, vp, vp, u64, vp, vp, vp, vp, vp]),
        "chq_fanout_device": (i32, [vp, vp, u32, u32, vp, vp, u64, vp, vp]),
        "chq_publish_flags_device": (i32, [vp, vp, vp, vp, u32, i32, vp, u32, vp, vp, vp]),
        "chq_fanout_last_error": (C.c_char_p, []),
        "chq_deliver_encode_device": (i32, [vp, vp, u64, vp, vp, u32, u32, vp, u64, vp, vp, vp]),
        "chq_encode_last_error": (C.c_char_p, []),
        "chq_numa_bind_to_device": (i32, [i32, C.POINTER(i32), C.POINTER(i32)]),
        "chq_host_alloc": (i32, [C.POINTER(vp), u64]),
        "chq_host_free": (i32, [vp]),
        "chq_stats": (i32, [vp, C.c_char_p, u32]),
        "chq_profile": (i32, [vp, i32, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(u32)]),
        "chq_profile_kernels": (i32, [vp, i32, C.POINTER(C.c_double), C.POINTER(u32)]),
        "chq_kernel_launches": (u64, []),
        "chq_debug_limits": (i32, [vp, u32, u32]),
        "chq_debug_check": (i32, [vp]),
        "chq_debug_clocks": (i32, [C.POINTER(C.c_uint64)]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    _lib = L
    return L


def check(rc):
    if rc != CHQ_OK:
        raise ChqError(rc, (load_library().chq_last_error() or b"").decode("utf-8", "replace"))
    return rc
