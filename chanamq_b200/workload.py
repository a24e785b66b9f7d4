"""ctypes binding of the synthetic workload generators (chanamq_b200/workload/gen.cpp)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_L = None


def _lib():
    global _L
    if _L is None:
        path = os.path.join(_HERE, "libchqgen.so")
        if not os.path.exists(path):
            from .build import build_gen
            build_gen()
        L = C.CDLL(path)
        L.chqgen_topic_new.restype = C.c_void_p
        L.chqgen_topic_new.argtypes = [C.c_uint64]
        L.chqgen_topic_free.argtypes = [C.c_void_p]
        L.chqgen_topic_max_len.restype = C.c_uint32
        L.chqgen_topic_keys.restype = C.c_uint64
        L.chqgen_topic_keys.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        L.chqgen_topic_patterns.restype = C.c_uint64
        L.chqgen_topic_patterns.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p]
        _L = L
    return _L


class TopicWorkload:
    """T(seed, P, N, mode) of SURVEY.md §8(d)."""

    def __init__(self, seed):
        self.L = _lib()
        self.g = self.L.chqgen_topic_new(seed)
        self.max_len = self.L.chqgen_topic_max_len()

    def __del__(self):
        if getattr(self, "g", None):
            self.L.chqgen_topic_free(self.g)
            self.g = None

    def keys(self, n, stream=0, out_blob=None, out_off=None):
        """-> (blob uint8[bytes], off uint32[n+1]); optional preallocated (e.g. pinned) outputs"""
        blob = out_blob if out_blob is not None else np.empty(n * self.max_len + 16, dtype=np.uint8)
        off = out_off if out_off is not None else np.empty(n + 1, dtype=np.uint32)
        used = self.L.chqgen_topic_keys(self.g, stream, n, blob.ctypes.data, off.ctypes.data)
        return blob[:max(int(used), 1)], off

    def patterns(self, p, n_queues=None, hash_tail=False):
        """-> (blob, off, queue_idx uint32[p], durable uint8[p]); queue names are queue_name(idx)"""
        nq = n_queues or p
        blob = np.empty(p * (self.max_len + 2) + 16, dtype=np.uint8)
        off = np.empty(p + 1, dtype=np.uint32)
        q = np.empty(p, dtype=np.uint32)
        d = np.empty(p, dtype=np.uint8)
        used = self.L.chqgen_topic_patterns(self.g, p, nq, int(hash_tail), blob.ctypes.data, off.ctypes.data,
                                            q.ctypes.data, d.ctypes.data)
        return blob[:max(int(used), 1)], off, q, d


def amqp_stream(topic, target_bytes, n_seg, body_bytes=4096, stream=0, out=None):
    """AMQP 0-9-1 publish stream of SURVEY.md §8(d) config 5 -> (bytes uint8[], seg_offsets uint64[n_seg+1], n_messages).
    `topic` is a TopicWorkload (routing keys come from it); `out` may be a preallocated (pinned) buffer."""
    L = _lib()
    L.chqgen_amqp_max_message.restype = C.c_uint64
    L.chqgen_amqp_max_message.argtypes = [C.c_uint32]
    L.chqgen_amqp_stream.restype = C.c_uint64
    L.chqgen_amqp_stream.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64,
                                     C.c_void_p, C.POINTER(C.c_uint64)]
    buf = out if out is not None else np.empty(int(target_bytes) + 64, dtype=np.uint8)
    seg = np.zeros(n_seg + 1, dtype=np.uint64)
    n_msgs = C.c_uint64(0)
    used = L.chqgen_amqp_stream(topic.g, stream, int(target_bytes), n_seg, body_bytes, buf.ctypes.data, int(buf.size),
                                seg.ctypes.data, C.byref(n_msgs))
    return buf[: int(used)], seg, int(n_msgs.value)


def queue_names_packed(queue_idx):
    """queue index -> name "q%07u", packed (blob, off); ascending index = ascending name"""
    n = len(queue_idx)
    digits = np.zeros((n, 8), dtype=np.uint8)
    digits[:, 0] = ord("q")
    v = queue_idx.astype(np.uint64).copy()
    for c in range(7, 0, -1):
        digits[:, c] = (v % 10).astype(np.uint8) + ord("0")
        v //= 10
    off = (np.arange(n + 1, dtype=np.uint64) * 8).astype(np.uint32)
    return digits.reshape(-1), off


class HeadersWorkload:
    """Config 3 of SURVEY.md §8(d): headers exchange, 8 fields per message, 8 argument slots per binding
    (chanamq_b200/workload/gen_headers.cpp)."""

    def __init__(self, seed, n_templates=100_000):
        L = self.L = _lib()
        L.chqgen_headers_new.restype = C.c_void_p
        L.chqgen_headers_new.argtypes = [C.c_uint64, C.c_uint32]
        L.chqgen_headers_free.argtypes = [C.c_void_p]
        L.chqgen_headers_max_bytes.restype = C.c_uint32
        L.chqgen_headers_messages.restype = C.c_uint64
        L.chqgen_headers_messages.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        L.chqgen_headers_bindings.restype = C.c_uint64
        L.chqgen_headers_bindings.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.g = L.chqgen_headers_new(seed, n_templates)
        self.max_bytes = L.chqgen_headers_max_bytes()

    def __del__(self):
        if getattr(self, "g", None):
            self.L.chqgen_headers_free(self.g)
            self.g = None

    def messages(self, n, stream=0):
        """-> (blob uint8[], off uint32[n+1]): the messages' headers tables (field-table contents)"""
        blob = np.empty(n * self.max_bytes + 16, dtype=np.uint8)
        off = np.empty(n + 1, dtype=np.uint32)
        used = self.L.chqgen_headers_messages(self.g, stream, n, blob.ctypes.data, off.ctypes.data)
        return blob[:max(int(used), 1)], off

    def bindings(self, p, n_queues=None):
        """-> (blob, off uint32[p+1], queue_idx uint32[p], durable uint8[p]): argument tables incl. x-match"""
        blob = np.empty(p * self.max_bytes + 16, dtype=np.uint8)
        off = np.empty(p + 1, dtype=np.uint32)
        q = np.empty(p, dtype=np.uint32)
        d = np.empty(p, dtype=np.uint8)
        used = self.L.chqgen_headers_bindings(self.g, p, n_queues or p, blob.ctypes.data, off.ctypes.data, q.ctypes.data,
                                              d.ctypes.data)
        return blob[:max(int(used), 1)], off, q, d
