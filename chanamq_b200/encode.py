"""Delivery-side encode (chana-mq-base/.../model/AMQCommand.scala:30-65): Basic.Deliver + content header +
body frames for a batch of deliveries, rendered on the device from the publish batch the decode kernels described.
Thin ctypes wrapper over chq_deliver_encode_device; arguments are raw device pointers (ints)."""
import numpy as np

from . import _lib
from ._lib import ChqError

DELIVERY_DTYPE = np.dtype([("delivery_tag", "<u8"), ("rec", "<u4"), ("ctag_off", "<u4"), ("channel", "<u2"), ("ctag_len", "u1"),
                           ("redelivered", "u1"), ("pad", "<u4")])
assert DELIVERY_DTYPE.itemsize == 24


def deliver_encode_device(d_bytes, d_recs, n_recs, d_ctags, d_deliveries, n_del, frame_max, d_out, out_cap, d_out_offsets, d_status,
                          stream=0):
    """asynchronous on `stream`; d_out_offsets = uint64[n_del + 1], d_status = int32"""
    L = _lib.load_library()
    rc = L.chq_deliver_encode_device(d_bytes, d_recs, n_recs, d_ctags, d_deliveries, n_del, frame_max, d_out, out_cap, d_out_offsets,
                                     d_status, stream)
    if rc != 0:
        raise ChqError(rc, (L.chq_encode_last_error() or b"").decode("utf-8", "replace"))
