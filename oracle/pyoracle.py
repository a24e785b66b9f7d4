"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; nothing under chanamq_b200/ does.  See oracle/trie_compat.hpp for what it restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DIRECT, FANOUT, TOPIC, HEADERS = 0, 1, 2, 3
COMPAT, SPEC = 0, 1


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        u8p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)
        L.orc_new.restype = C.c_void_p
        L.orc_new.argtypes = [C.c_int, C.c_int]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_set_inplace.argtypes = [C.c_void_p, C.c_int]
        for f in (L.orc_bind, L.orc_unbind):
            f.restype = C.c_int
            f.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_int]
        L.orc_bind_many.restype = C.c_int
        L.orc_bind_many.argtypes = [C.c_void_p] * 6 + [C.c_uint32]
        L.orc_n_queues.restype = C.c_uint32
        L.orc_n_queues.argtypes = [C.c_void_p]
        L.orc_queue_name.restype = C.c_int
        L.orc_queue_name.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), u32p]
        L.orc_route.restype = C.c_int64
        L.orc_route.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int]
        L.orc_offsets.restype = C.c_void_p
        L.orc_ids.restype = C.c_void_p
        L.orc_durable.restype = C.c_void_p
        L.orc_err.restype = C.c_void_p
        for f in (L.orc_offsets, L.orc_ids, L.orc_durable, L.orc_err):
            f.argtypes = [C.c_void_p]
        L.orc_last_seconds.restype = C.c_double
        L.orc_last_seconds.argtypes = [C.c_void_p]
        L.orc_pair_match.restype = C.c_int
        L.orc_pair_match.argtypes = [C.c_int, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32]
        L.orc_split.restype = C.c_int
        L.orc_split.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, u32p]
        _LIB = L
    return _LIB


def _b(s):
    return s if isinstance(s, (bytes, bytearray)) else s.encode("utf-8")


def pack_keys(keys):
    """list of str/bytes -> (uint8 blob, uint32 offsets[n+1])"""
    bs = [_b(k) for k in keys]
    off = np.zeros(len(bs) + 1, dtype=np.uint32)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64).astype(np.uint32)
    blob = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
    return blob, off


def split(s):
    s = _b(s)
    out = C.create_string_buffer(len(s) + 1)
    n = C.c_uint32(0)
    cnt = lib().orc_split(s, len(s), out, len(s) + 1, C.byref(n))
    if cnt <= 0:
        return []
    return out.raw[: n.value].split(b"\n")


def pair_match(mode, pattern, key):
    p, k = _b(pattern), _b(key)
    return lib().orc_pair_match(mode, p, len(p), k, len(k))


REC_DTYPE = np.dtype([  # mirrors chq_publish_rec (oracle/amqp_decode.cpp Rec)
    ("method_off", "<u8"), ("body_off", "<u8"), ("body_size", "<u8"), ("ttl", "<i8"),
    ("exch_off", "<u4"), ("exch_len", "<u4"), ("key_off", "<u4"), ("key_len", "<u4"),
    ("hdr_tbl_off", "<u4"), ("hdr_tbl_len", "<u4"), ("seg", "<u4"), ("channel", "<u2"), ("kind", "u1"), ("flags", "u1"),
])


def decode(data, seg_offsets, chunk=0):
    """Restated FrameParser + CommandAssembler + property readers over n_seg connection segments, each fed to
    its own parser in chunks of `chunk` bytes (0 = all at once) -> (recs, seg_consumed, seg_err, seg_err_off)."""
    L = lib()
    L.orc_decode.restype = C.c_int
    L.orc_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_uint64,
                             C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p, C.c_void_p]
    data = np.ascontiguousarray(data, dtype=np.uint8)
    seg = np.ascontiguousarray(seg_offsets, dtype=np.uint64)
    n_seg = len(seg) - 1
    cap = len(data) // 12 + n_seg + 1
    recs = np.zeros(cap, dtype=REC_DTYPE)
    n = C.c_uint64(0)
    consumed = np.zeros(max(n_seg, 1), dtype=np.uint64)
    err = np.zeros(max(n_seg, 1), dtype=np.int32)
    err_off = np.zeros(max(n_seg, 1), dtype=np.uint64)
    buf = data if data.size else np.zeros(1, dtype=np.uint8)
    rc = L.orc_decode(buf.ctypes.data, seg.ctypes.data, n_seg, chunk, recs.ctypes.data, cap, C.byref(n), consumed.ctypes.data,
                      err.ctypes.data, err_off.ctypes.data)
    if rc != 0:
        raise RuntimeError("oracle decode: record capacity")
    return recs[: n.value].copy(), consumed[:n_seg], err[:n_seg], err_off[:n_seg]


class HeadersOracle:
    """Brute-force x-match all/any over (message, binding) pairs (oracle/headers_spec.cpp; parity unpinned)."""

    def __init__(self):
        L = self.L = lib()
        L.orch_new.restype = C.c_void_p
        L.orch_free.argtypes = [C.c_void_p]
        L.orch_bind.restype = C.c_int
        L.orch_bind.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_int, C.c_char_p, C.c_uint32]
        L.orch_n_queues.restype = C.c_uint32
        L.orch_n_queues.argtypes = [C.c_void_p]
        L.orch_queue_name.restype = C.c_int
        L.orch_queue_name.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
        L.orch_route.restype = C.c_int64
        L.orch_route.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        for f in (L.orch_offsets, L.orch_ids, L.orch_durable, L.orch_err):
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p]
        self.h = L.orch_new()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orch_free(self.h)
            self.h = None

    def bind(self, queue, durable, args_table):
        q, a = _b(queue), bytes(args_table)
        return self.L.orch_bind(self.h, q, len(q), int(durable), a, len(a))

    def queue_names(self):
        out = []
        p, n = C.c_void_p(), C.c_uint32()
        for i in range(self.L.orch_n_queues(self.h)):
            self.L.orch_queue_name(self.h, i, C.byref(p), C.byref(n))
            out.append(C.string_at(p, n.value))
        return out

    def route_packed(self, blob, off):
        n = len(off) - 1
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        total = self.L.orch_route(self.h, blob.ctypes.data, off.ctypes.data, n)

        def arr(ptr, count, dt):
            if count == 0:
                return np.zeros(0, dtype=dt)
            buf = (C.c_uint8 * (count * np.dtype(dt).itemsize)).from_address(ptr)
            return np.frombuffer(buf, dtype=dt, count=count).copy()

        return (arr(self.L.orch_offsets(self.h), n + 1, np.uint32), arr(self.L.orch_ids(self.h), int(total), np.uint32),
                arr(self.L.orch_durable(self.h), n, np.uint8), arr(self.L.orch_err(self.h), n, np.uint8))


class Oracle:
    """One exchange's matcher, restated from the reference (trait QueueMatcher, QueueMatcher.scala:11-27)."""

    def __init__(self, exchange_type=TOPIC, mode=COMPAT, inplace=False):
        self.L = lib()
        self.h = self.L.orc_new(exchange_type, mode)
        if inplace:
            self.L.orc_set_inplace(self.h, 1)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_free(self.h)
            self.h = None

    def bind(self, key, queue, durable=False):
        k, q = _b(key), _b(queue)
        return self.L.orc_bind(self.h, k, len(k), q, len(q), int(durable))

    def unbind(self, key, queue, durable=False):
        k, q = _b(key), _b(queue)
        return self.L.orc_unbind(self.h, k, len(k), q, len(q), int(durable))

    def bind_many(self, keys_blob, key_off, queues_blob, queue_off, durable):
        a = [np.ascontiguousarray(x) for x in (keys_blob, key_off, queues_blob, queue_off, durable)]
        return self.L.orc_bind_many(self.h, *[x.ctypes.data for x in a], len(key_off) - 1)

    def n_queues(self):
        return self.L.orc_n_queues(self.h)

    def queue_name(self, qid):
        p, n = C.c_void_p(), C.c_uint32()
        if self.L.orc_queue_name(self.h, qid, C.byref(p), C.byref(n)) != 0:
            raise IndexError(qid)
        return C.string_at(p, n.value)

    def queue_names(self):
        return [self.queue_name(i) for i in range(self.n_queues())]

    def route_packed(self, blob, off, threads=1):
        """-> (offsets u32[n+1], ids u32[total], any_durable u8[n], err u8[n]) as numpy copies"""
        n = len(off) - 1
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        total = self.L.orc_route(self.h, blob.ctypes.data, off.ctypes.data, n, threads)
        if total < 0:
            raise RuntimeError("oracle route overflow")

        def arr(ptr, count, dt):
            if count == 0:
                return np.zeros(0, dtype=dt)
            buf = (C.c_uint8 * (count * np.dtype(dt).itemsize)).from_address(ptr)
            return np.frombuffer(buf, dtype=dt, count=count).copy()

        return (
            arr(self.L.orc_offsets(self.h), n + 1, np.uint32),
            arr(self.L.orc_ids(self.h), int(total), np.uint32),
            arr(self.L.orc_durable(self.h), n, np.uint8),
            arr(self.L.orc_err(self.h), n, np.uint8),
        )

    def last_seconds(self):
        return self.L.orc_last_seconds(self.h)

    def route(self, keys, threads=1):
        """-> list (per key) of sorted queue-name lists; None where the reference would throw (zero words)"""
        blob, off = pack_keys(keys)
        o, ids, _, err = self.route_packed(blob, off, threads)
        names = self.queue_names()
        return [None if err[i] else [names[j] for j in ids[o[i]:o[i + 1]]] for i in range(len(keys))]


# ---- the step after the match (test infrastructure; chana-mq-server/.../entity/ExchangeEntity.scala:300-307, 319-330) ----
def fanout_transpose(offsets, queue_ids, n_queues):
    """CSR message -> queue ids  =>  (q_offsets[n_queues + 1], q_msgs[total]): every queue's messages in publish order,
    which is what `queues foreach { q => queueToMsgs += (q -> (queueToMsgs(q) :+ Msg(...))) }` builds message by message"""
    off = np.asarray(offsets, dtype=np.int64)
    ids = np.asarray(queue_ids, dtype=np.int64)
    msg = np.repeat(np.arange(len(off) - 1, dtype=np.int64), np.diff(off))
    order = np.argsort(ids, kind="stable")
    q_off = np.zeros(n_queues + 1, dtype=np.uint32)
    q_off[1:] = np.cumsum(np.bincount(ids, minlength=n_queues))
    return q_off, msg[order].astype(np.uint32)


def publish_flags(offsets, queue_ids, any_durable, pub_flags, exchange_durable, queue_has_consumer=None):
    """-> (refer_count uint32[n], flags uint8[n]): bit 0 isPersistent (EX:302), bit 1 isNonRouted (mandatory && no
    queue), bit 2 isNonDeliverable (immediate && no matched queue has a consumer) (EX:319-330)"""
    off = np.asarray(offsets, dtype=np.int64)
    n = len(off) - 1
    cnt = np.diff(off)
    pf = np.zeros(n, dtype=np.uint8) if pub_flags is None else np.asarray(pub_flags, dtype=np.uint8)
    out = np.zeros(n, dtype=np.uint8)
    out |= (bool(exchange_durable) & ((pf & 4) != 0) & (np.asarray(any_durable) != 0)).astype(np.uint8)
    out |= ((((pf & 1) != 0) & (cnt == 0)).astype(np.uint8) << 1)
    has = np.zeros(n, dtype=bool)
    if queue_has_consumer is not None and len(queue_ids):
        hc = np.asarray(queue_has_consumer)[np.asarray(queue_ids, dtype=np.int64)] != 0
        msg = np.repeat(np.arange(n), cnt)
        has = np.bincount(msg, weights=hc, minlength=n) > 0
    out |= ((((pf & 2) != 0) & ~has).astype(np.uint8) << 2)
    return cnt.astype(np.uint32), out


# ---- delivery-side encode (test infrastructure; small cases only: plain Python) ----
def render_deliver(channel, consumer_tag, delivery_tag, redelivered, exchange, routing_key, header_payload, body, frame_max):
    """AMQCommand.render for a Basic.Deliver (chana-mq-base/.../model/AMQCommand.scala:39-65): METHOD frame with the
    arguments of .../method/Basic.scala:198-204, the HEADER frame AMQContentHeader.toFrame builds (class id, weight 0,
    bodySize = body.length, then the properties: `header_payload` is that payload as the restated property writer of
    tests/amqp_build.py produced it; only its weight and bodySize are overwritten here), BODY frames of at most
    frameMax - NON_BODY_SIZE payload bytes (one frame of everything if frameMax == 0), every frame as Frame.writeTo does
    (.../model/Frame.scala:220-226)."""
    import struct

    def frame(t, payload):
        return struct.pack(">BHI", t, channel, len(payload)) + payload + b"\xce"

    def shortstr(s):
        return bytes([len(s)]) + s

    args = (struct.pack(">HH", 60, 60) + shortstr(_b(consumer_tag)) + struct.pack(">Q", delivery_tag) + bytes([1 if redelivered else 0]) +
            shortstr(_b(exchange)) + shortstr(_b(routing_key)))
    out = frame(1, args)
    hp = bytes(header_payload)
    out += frame(2, hp[:2] + struct.pack(">HQ", 0, len(body)) + hp[12:])
    bmax = len(body) if frame_max == 0 else frame_max - 8
    off = 0
    while off < len(body):
        out += frame(3, body[off:off + bmax])
        off += bmax
    return out
