"""The step after the match (chana-mq-server/.../entity/ExchangeEntity.scala:300-307, 319-330):
per-queue message lists in publish order and the per-message flags, computed on the device from a routed batch.
Thin ctypes wrappers over chq_fanout_device / chq_publish_flags_device; arguments are raw device pointers (ints)."""
from . import _lib
from ._lib import ChqError

FAN_PERSISTENT, FAN_NON_ROUTED, FAN_NON_DELIVERABLE = 1, 2, 4


def _check(rc):
    if rc != 0:
        L = _lib.load_library()
        raise ChqError(rc, (L.chq_fanout_last_error() or b"").decode("utf-8", "replace"))


def fanout_device(d_offsets, d_queue_ids, n, n_queues, d_q_offsets, d_q_msgs, q_cap, d_status, stream=0):
    """CSR message -> queue ids  =>  CSR queue -> message indices (ascending); asynchronous on `stream`"""
    _check(_lib.load_library().chq_fanout_device(d_offsets, d_queue_ids, n, n_queues, d_q_offsets, d_q_msgs, q_cap, d_status, stream))


def publish_flags_device(d_offsets, d_queue_ids, d_any_durable, d_pub_flags, n, exchange_durable, d_queue_has_consumer, n_queues,
                         d_refer_count, d_out_flags, stream=0):
    _check(_lib.load_library().chq_publish_flags_device(d_offsets, d_queue_ids, d_any_durable, d_pub_flags, n, int(exchange_durable),
                                                        d_queue_has_consumer, n_queues, d_refer_count, d_out_flags, stream))
