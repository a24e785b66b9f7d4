"""chanamq_b200 — B200-native batched exchange routing for ChanaMQ behind the reference's
QueueMatcher seam (chana-mq-server/.../engine/QueueMatcher.scala:11-27).

The product is the C-ABI library libchanagpu.so (include/chanamq_gpu.h); this package is the
Python mirror of the reference-side interface used by the tests and the benchmark.  There is no
CPU fallback: importing works anywhere, but creating a table needs a CUDA device.
"""
from ._lib import (CHQ_DIRECT, CHQ_FANOUT, CHQ_TOPIC, CHQ_HEADERS, CHQ_MODE_COMPAT, CHQ_MODE_SPEC, ChqError,
                   load_library)
from .matcher import GpuMatcher, Subscriber, Subscription, RouteResult

__all__ = ["GpuMatcher", "Subscriber", "Subscription", "RouteResult", "ChqError", "load_library", "CHQ_DIRECT",
           "CHQ_FANOUT", "CHQ_TOPIC", "CHQ_HEADERS", "CHQ_MODE_COMPAT", "CHQ_MODE_SPEC"]
