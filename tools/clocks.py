"""Cycle accounting of k_write_rest (the messages the main write kernel leaves out) by kind of message.  Needs a library built with CHQ_NVCC_DEFS=-DCHQ_CLOCKS:
   CHQ_NVCC_DEFS=-DCHQ_CLOCKS python -c "import __graft_entry__ as g; g.build()"; python tools/clocks.py [compat|spec]"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chanamq_b200 import GpuMatcher, CHQ_TOPIC, CHQ_MODE_COMPAT, CHQ_MODE_SPEC
from chanamq_b200.workload import TopicWorkload, queue_names_packed

mode = CHQ_MODE_SPEC if (len(sys.argv) > 1 and sys.argv[1] == "spec") else CHQ_MODE_COMPAT
w = TopicWorkload(4)
pb, po, qi, dur = w.patterns(1_000_000, hash_tail=True)[:4]
qb, qo = queue_names_packed(qi)
m = GpuMatcher(CHQ_TOPIC, mode)
m.bind_many(pb, po, qb, qo, dur)
m.commit()
blob, off = w.keys(1 << 20, stream=0)
L = m.L
out = (C.c_uint64 * 32)()
for it in range(3):
    r = m.route_packed(blob, off, copy=False)
    L.chq_debug_clocks(out)
v = list(out)
names = ["sweep for big", "tiny", "single list", "slow-path copy", "multi-way", "weave <=32 short", "weave >32 short", "chunk header"]
tot = sum(v[:8]) or 1
print("mode", "spec" if mode == CHQ_MODE_SPEC else "compat", "ids", len(r.queue_ids))
for b, nm in enumerate(names):
    print("%-18s %9.1f M warp-cycles %5.1f%%  msgs %8d  ids %10d  cycles/msg %8.0f" % (
        nm, v[b] / 1e6, 100.0 * v[b] / tot, v[8 + b], v[16 + b], v[b] / max(1, v[8 + b])))
print("total %.1f M warp-cycles" % (tot / 1e6))
