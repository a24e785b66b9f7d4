"""One-GPU look at the merge stage of binding-slice sharding (chq_shard_route_device step 4) at benchmark scale: G slice
tables of the 1 M-binding exchange on one device, one batch routed against each, the slices' ids of the first n / G
messages laid out as k_pull leaves them, then chq_merge_slices_device2 timed alone (CUDA events), so that the kernels
can be profiled with ncu without a multi-rank run.   python tools/bench_merge.py [G] [steps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_TOPIC, GpuMatcher
from chanamq_b200._lib import check, load_library
from chanamq_b200.sharded import slice_bindings
from chanamq_b200.workload import TopicWorkload, queue_names_packed

G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
L = load_library()
w = TopicWorkload(4)
pb, poff, q, d = w.patterns(1_000_000, 1_000_000, hash_tail=True)
qb, qoff = queue_names_packed(q)
binds = (pb, poff, qb, qoff, d)
n = 1 << 20
blob, off = w.keys(n, stream=5000)
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
d_blob = torch.from_numpy(blob.copy()).to(dev)
d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
cap = int(n * (100 / G * 1.3 + 8))
nn = n // G
offs, ids, durs = [], [], []
for r in range(G):
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    m.bind_many(*slice_bindings(*binds, r, G))
    check(L.chq_declare_queues(m.h, qb.ctypes.data, qoff.ctypes.data, len(qoff) - 1))
    m.commit()
    o = torch.empty(n + 1, dtype=torch.int32, device=dev)
    i = torch.empty(cap, dtype=torch.int32, device=dev)
    du = torch.empty(n, dtype=torch.uint8, device=dev)
    er = torch.empty(n, dtype=torch.uint8, device=dev)
    tot = torch.zeros(2, dtype=torch.int64, device=dev)
    m.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, o.data_ptr(), i.data_ptr(), cap, du.data_ptr(), er.data_ptr(),
                   tot.data_ptr(), tot.data_ptr() + 8, st)
    torch.cuda.synchronize()
    assert int(tot[1]) & 0xFFFFFFFF == 0
    e = int(o[nn])
    offs.append(o[: nn + 1].clone())
    ids.append(i[:e].clone())
    durs.append(du[:nn].clone())
    del m, i
pull_cap = max(len(x) for x in ids) + 64
got = torch.zeros(G * pull_cap, dtype=torch.int32, device=dev)
for r in range(G):
    got[r * pull_cap: r * pull_cap + len(ids[r])] = ids[r]
sub_off = torch.cat(offs).contiguous()
sub_dur = torch.cat(durs).contiguous()
out_cap = sum(len(x) for x in ids) + 1024
out_off = torch.empty(nn + 1, dtype=torch.int32, device=dev)
out_ids = torch.empty(out_cap, dtype=torch.int32, device=dev)
out_dur = torch.empty(nn, dtype=torch.uint8, device=dev)
tot = torch.zeros(2, dtype=torch.int64, device=dev)


def merge():
    check(L.chq_merge_slices_device2(0, sub_off.data_ptr(), got.data_ptr(), pull_cap, sub_dur.data_ptr(), G, nn, out_off.data_ptr(),
                                     out_ids.data_ptr(), out_cap, out_dur.data_ptr(), tot.data_ptr(), tot.data_ptr() + 8, 1, st))


for _ in range(3):
    merge()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    merge()
e1.record()
torch.cuda.synchronize()
print("G=%d: merge of %d messages, %d ids: %.3f ms per batch, status %d" % (
    G, nn, int(tot[0]), e0.elapsed_time(e1) / steps, int(tot[1]) & 0xFFFFFFFF))
c = (out_off[1:] - out_off[:-1]).to(torch.int64).cpu().numpy()
print("ids per message: mean %.1f  p99 %d  max %d;  messages over 1024 ids: %d (%.1f %% of the ids), over 4096: %d (%.1f %%)" % (
    c.mean(), int(np.percentile(c, 99)), int(c.max()), int((c > 1024).sum()), 100.0 * c[c > 1024].sum() / max(1, c.sum()),
    int((c > 4096).sum()), 100.0 * c[c > 4096].sum() / max(1, c.sum())))
