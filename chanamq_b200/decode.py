"""Python mirror of the decode seam over the C ABI (chq_decode* in include/chanamq_gpu.h): what
FrameStage gets from FrameParser + CommandAssembler + the Basic.Publish / content-header readers
(chana-mq-server/.../engine/FrameStage.scala:291-331, 483-508), for many connection segments at once."""
import ctypes as C
from typing import NamedTuple

import numpy as np

from . import _lib
from ._lib import check

# chq_publish_rec, 64 bytes
REC_DTYPE = np.dtype([
    ("method_off", "<u8"), ("body_off", "<u8"), ("body_size", "<u8"), ("ttl", "<i8"),
    ("exch_off", "<u4"), ("exch_len", "<u4"), ("key_off", "<u4"), ("key_len", "<u4"),
    ("hdr_tbl_off", "<u4"), ("hdr_tbl_len", "<u4"), ("seg", "<u4"), ("channel", "<u2"), ("kind", "u1"), ("flags", "u1"),
])
assert REC_DTYPE.itemsize == 64
PUB_MANDATORY, PUB_IMMEDIATE, PUB_PERSIST, PUB_HAS_TTL, PUB_HAS_HEADERS, PUB_HAS_PROPS = 1, 2, 4, 8, 16, 32


class DecodeResult(NamedTuple):
    recs: np.ndarray          # REC_DTYPE[n_recs], segment by segment in stream order
    seg_consumed: np.ndarray  # uint64[n_seg]
    seg_err: np.ndarray       # int32[n_seg]: 0, 501 (FRAME_ERROR), 1 (the reference throws)
    seg_err_off: np.ndarray   # uint64[n_seg]


def decode(data, seg_offsets, recs_cap=None):
    """host bytes in, host records out (chq_decode).  Needs a CUDA device: there is no CPU decoder here."""
    L = _lib.load_library()
    data = np.ascontiguousarray(data, dtype=np.uint8)
    seg = np.ascontiguousarray(seg_offsets, dtype=np.uint64)
    n_seg = len(seg) - 1
    cap = int(recs_cap if recs_cap is not None else len(data) // 12 + n_seg + 1)
    recs = np.zeros(cap, dtype=REC_DTYPE)
    n = C.c_uint64(0)
    consumed = np.zeros(max(n_seg, 1), dtype=np.uint64)
    err = np.zeros(max(n_seg, 1), dtype=np.int32)
    err_off = np.zeros(max(n_seg, 1), dtype=np.uint64)
    buf = data if data.size else np.zeros(1, dtype=np.uint8)
    check(L.chq_decode(buf.ctypes.data, int(data.size), seg.ctypes.data, n_seg, recs.ctypes.data, cap, C.byref(n),
                       consumed.ctypes.data, err.ctypes.data, err_off.ctypes.data))
    return DecodeResult(recs[: n.value].copy(), consumed[:n_seg], err[:n_seg], err_off[:n_seg])


def decode_device(d_bytes, length, d_seg_offsets, n_seg, d_recs, recs_cap, d_n_recs, d_consumed, d_err, d_err_off, d_status,
                  stream=0):
    """raw device pointers, asynchronous on `stream` (chq_decode_device)"""
    check(_lib.load_library().chq_decode_device(d_bytes, length, d_seg_offsets, n_seg, d_recs, recs_cap, d_n_recs, d_consumed,
                                                d_err, d_err_off, d_status, stream))


def publish_keys_device(d_recs, d_n_recs, recs_cap, d_key_off, d_key_len, d_rec_index, d_n_pub, stream=0):
    check(_lib.load_library().chq_publish_keys_device(d_recs, d_n_recs, recs_cap, d_key_off, d_key_len, d_rec_index, d_n_pub,
                                                      stream))
