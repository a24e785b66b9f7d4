"""Python mirror of the reference's matcher interface over the C ABI.

  trait QueueMatcher { subscribe(topic, sub): Subscription; unsubscribe(sub); lookup(topic): Set[Subscriber] }
      — chana-mq-server/.../engine/QueueMatcher.scala:11-27
  TopicSubscriber(queue, isDurable)  — .../entity/ExchangeEntity.scala:97

plus the one ADDED batch call the exchange actor makes instead of its per-message loop
(ExchangeEntity.scala:293-308): lookup_all / route_packed.
"""
import ctypes as C
import json
from typing import NamedTuple

import numpy as np

from . import _lib
from ._lib import ChqResult, check


class Subscriber(NamedTuple):  # ExchangeEntity.TopicSubscriber
    queue: str
    durable: bool = False


class Subscription(NamedTuple):  # QueueMatcher.scala:9
    topic: str
    subscriber: Subscriber


class RouteResult(NamedTuple):
    offsets: np.ndarray      # uint32[n+1]
    queue_ids: np.ndarray    # uint32[offsets[n]], ascending within a message
    any_durable: np.ndarray  # uint8[n]
    err: np.ndarray          # uint8[n]: 1 zero-word key (the reference throws), 2 key longer than 255 bytes


def _b(s):
    return s if isinstance(s, (bytes, bytearray)) else s.encode("utf-8")


def pack(strings):
    bs = [_b(k) for k in strings]
    off = np.zeros(len(bs) + 1, dtype=np.uint32)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64).astype(np.uint32)
    blob = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
    if blob.size == 0:
        blob = np.zeros(1, dtype=np.uint8)
    return blob, off


def _view(ptr, count, dtype):
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    nbytes = count * np.dtype(dtype).itemsize
    return np.frombuffer((C.c_uint8 * nbytes).from_address(ptr), dtype=dtype, count=count)


class GpuMatcher:
    """One exchange's matcher on one B200.  Binds take effect at the next commit() (implicit before a lookup)."""

    def __init__(self, exchange_type=_lib.CHQ_TOPIC, mode=_lib.CHQ_MODE_COMPAT, device=0):
        self.L = _lib.load_library()
        h = C.c_void_p()
        check(self.L.chq_table_create(C.byref(h), exchange_type, mode, device))
        self.h = h
        self._dirty = True
        self._names = None

    def close(self):
        if getattr(self, "h", None):
            self.L.chq_table_destroy(self.h)
            self.h = None

    __del__ = close

    # ---- QueueMatcher ----------------------------------------------------------------------
    def subscribe(self, topic, sub):
        sub = sub if isinstance(sub, Subscriber) else Subscriber(*sub)
        k, q = _b(topic), _b(sub.queue)
        check(self.L.chq_bind(self.h, k, len(k), q, len(q), int(sub.durable)))
        self._dirty = True
        return Subscription(topic, sub)

    def unsubscribe(self, subscription):
        k, q = _b(subscription.topic), _b(subscription.subscriber.queue)
        check(self.L.chq_unbind(self.h, k, len(k), q, len(q), int(subscription.subscriber.durable)))
        self._dirty = True

    def lookup(self, topic):
        """Set of queue names for one key (the slow, API-compatible path: a batch of one)."""
        r = self.lookup_all([topic])[0]
        if r is None:
            raise IndexError("zero-word topic")  # ArrayIndexOutOfBoundsException in the reference
        return set(r)

    # ---- table management -------------------------------------------------------------------
    def bind(self, key, queue, durable=False):
        return self.subscribe(key, Subscriber(queue if isinstance(queue, str) else queue.decode(), durable))

    def try_bind(self, key, queue, durable=False):
        """status code instead of an exception (0, CHQ_E_ZERO_WORDS, ...)"""
        k, q = _b(key), _b(queue)
        rc = self.L.chq_bind(self.h, k, len(k), q, len(q), int(durable))
        self._dirty = True
        return rc

    def bind_args(self, key, queue, durable, args_table):
        """bind with the raw contents of the binding's arguments field table (headers exchanges)"""
        k, q, a = _b(key), _b(queue), bytes(args_table)
        rc = self.L.chq_bind_args(self.h, k, len(k), q, len(q), int(durable), a, len(a))
        self._dirty = True
        return rc

    def route_headers_packed(self, blob, off, copy=True):
        """headers exchange: blob/off = the messages' headers tables (field table contents), packed"""
        self._ensure()
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        res = ChqResult()
        check(self.L.chq_route(self.h, None, None, len(off) - 1, blob.ctypes.data, off.ctypes.data, C.byref(res)))
        return self._result(res, copy)

    def unbind(self, key, queue, durable=False):
        k, q = _b(key), _b(queue)
        check(self.L.chq_unbind(self.h, k, len(k), q, len(q), int(durable)))
        self._dirty = True

    def unbind_queue(self, queue):
        q = _b(queue)
        check(self.L.chq_unbind_queue(self.h, q, len(q)))
        self._dirty = True

    def bind_many(self, keys_blob, key_off, queues_blob, queue_off, durable):
        a = [np.ascontiguousarray(x) for x in (keys_blob, key_off, queues_blob, queue_off, durable)]
        check(self.L.chq_bind_many(self.h, *[x.ctypes.data for x in a], len(key_off) - 1))
        self._dirty = True

    def commit(self):
        """in place where the changes allow it, else a full compile (chq_commit)"""
        check(self.L.chq_commit(self.h))
        self._dirty = False
        self._names = None

    def commit_full(self):
        """compile the whole binding set again: queue ids become name ranks again (chq_commit_full)"""
        check(self.L.chq_commit_full(self.h))
        self._dirty = False
        self._names = None

    def _ensure(self):
        if self._dirty:
            self.commit()

    def n_queues(self):
        self._ensure()
        n = C.c_uint32()
        check(self.L.chq_n_queues(self.h, C.byref(n)))
        return n.value

    def queue_names(self):
        self._ensure()
        if self._names is None:
            out = []
            p, ln = C.c_void_p(), C.c_uint32()
            for i in range(self.n_queues()):
                check(self.L.chq_queue_name(self.h, i, C.byref(p), C.byref(ln)))
                out.append(C.string_at(p, ln.value))
            self._names = out
        return self._names

    def stats(self):
        self._ensure()
        buf = C.create_string_buffer(2048)
        check(self.L.chq_stats(self.h, buf, 2048))
        return json.loads(buf.value.decode())

    def profile(self, enable):
        """-> (tokenize_ms, route_ms, batches) summed since the last call; then switch recording on/off"""
        a, b, n = C.c_double(), C.c_double(), C.c_uint32()
        check(self.L.chq_profile(self.h, int(enable), C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value

    def profile_kernels(self, enable):
        """-> ({kernel: ms summed since the last call}, batches); then switch recording on/off"""
        ms, n = (C.c_double * 5)(), C.c_uint32()
        check(self.L.chq_profile_kernels(self.h, int(enable), ms, C.byref(n)))
        return dict(zip(("k_tokenize", "k_probe", "k_offsets", "k_write", "k_write_rest"), list(ms))), n.value

    def debug_limits(self, max_slots=0, max_small=0):
        check(self.L.chq_debug_limits(self.h, max_slots, max_small))

    # ---- batched routing --------------------------------------------------------------------
    def _result(self, res, copy=True):
        n = res.n
        off = _view(res.offsets, n + 1, np.uint32)
        total = int(off[n]) if n else 0
        if res.ids_format == _lib.CHQ_IDS_U24:  # packed 3-byte ids (chq_set_ids_format): uint8[3 * total]
            ids = _view(res.queue_ids, 3 * total, np.uint8)
        else:
            ids = _view(res.queue_ids, total, np.uint32)
        out = RouteResult(off, ids, _view(res.any_durable, n, np.uint8), _view(res.err, n, np.uint8))
        return RouteResult(*[a.copy() for a in out]) if copy else out

    def set_ids_format(self, fmt):
        """CHQ_IDS_U24: host results carry the queue ids as 3-byte little-endian values (RouteResult.queue_ids is
        then uint8[3 * total]; unpack_u24 expands it)."""
        check(self.L.chq_set_ids_format(self.h, fmt))

    def unpack_u24(self, packed):
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        out = np.empty(len(packed) // 3, dtype=np.uint32)
        self.L.chq_unpack_u24(packed.ctypes.data, len(out), out.ctypes.data)
        return out

    def route_packed(self, blob, off, copy=True):
        """blob uint8[], off uint32[n+1] (host) -> RouteResult (views into library memory unless copy)"""
        self._ensure()
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        res = ChqResult()
        check(self.L.chq_route(self.h, blob.ctypes.data, off.ctypes.data, len(off) - 1, None, None, C.byref(res)))
        return self._result(res, copy)

    def submit(self, slot, blob_ptr, off_ptr, n):
        self._ensure()
        check(self.L.chq_route_submit(self.h, slot, blob_ptr, off_ptr, n))

    def wait(self, slot, copy=False):
        res = ChqResult()
        check(self.L.chq_route_wait(self.h, slot, C.byref(res)))
        return self._result(res, copy)

    def route_keys_device(self, d_bytes, d_key_off, d_key_len, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total,
                          d_status, stream=0):
        """chq_route_keys_device: keys scattered inside a device byte batch (as the frame decoder leaves them)"""
        self._ensure()
        check(self.L.chq_route_keys_device(self.h, d_bytes, d_key_off, d_key_len, n, d_offsets, d_ids, ids_cap, d_durable,
                                           d_err, d_total, d_status, stream))

    def route_device_static(self, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total, d_status,
                            stream=0):
        """chq_route_device_static: same arguments and results as route_device, but the call allocates
        nothing (fixed in-library scratch; at most CHQ_STATIC_MAX_KEYS keys; calls must not overlap)"""
        self._ensure()
        check(self.L.chq_route_device_static(self.h, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err,
                                             d_total, d_status, stream))

    def lookup_all(self, keys):
        """list of keys -> per key the sorted list of queue names (bytes), None where the reference throws"""
        blob, off = pack(keys)
        r = self.route_packed(blob, off)
        names = self.queue_names()
        return [None if r.err[i] else [names[j] for j in r.queue_ids[r.offsets[i]:r.offsets[i + 1]]]
                for i in range(len(keys))]

    def route_device(self, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total, d_status, stream=0):
        """raw device pointers (ints); asynchronous on `stream` (a cudaStream_t handle)"""
        self._ensure()
        check(self.L.chq_route_device(self.h, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total,
                                      d_status, stream))
