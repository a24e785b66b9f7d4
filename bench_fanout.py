#!/usr/bin/env python
"""Secondary benchmark: the step after the match — per-queue message lists in publish order and
the per-message flags of ExchangeEntity.publish (chana-mq-server/.../entity/ExchangeEntity.scala:300-307, 319-330) —
on the headline workload (topic exchange, 1 M bindings, batches of 2^20 keys, COMPAT semantics).

  python bench_fanout.py [--bindings P] [--batch N] [--steps K] [--warmup W]

Prints one JSON line: messages/s of chq_fanout_device + chq_publish_flags_device on a routed batch that is
already in HBM (CUDA events), the algorithmic bytes (ids read twice — count and scatter —, message indices
written once and sorted in place, queue offsets), and the numpy oracle's rate on a sample.  bench.py stays the
headline; this line documents the first "next" row."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bindings", type=int, default=1_000_000)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-sample", type=int, default=1 << 17)
    a = ap.parse_args()
    import torch

    from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_TOPIC, GpuMatcher
    from chanamq_b200 import fanout as gf
    from chanamq_b200._lib import load_library
    from chanamq_b200.workload import TopicWorkload, queue_names_packed
    from oracle import pyoracle as po

    if not torch.cuda.is_available():
        raise SystemExit("bench_fanout.py: no CUDA device; the product has no CPU path")
    L = load_library()
    dev = torch.device("cuda:0")
    w = TopicWorkload(4)
    pb, poff, q, d = w.patterns(a.bindings, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    m.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    n, nq = a.batch, m.n_queues()
    blob, off = w.keys(n, stream=0)
    st = torch.cuda.current_stream().cuda_stream
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_koff = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    cap = 160 * n
    d_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_ids = torch.empty(cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    d_err = torch.empty(n, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    m.route_device(d_blob.data_ptr(), d_koff.data_ptr(), n, d_off.data_ptr(), d_ids.data_ptr(), cap, d_dur.data_ptr(),
                   d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
    torch.cuda.synchronize()
    total = int(d_tot[0])
    assert (int(d_tot[1]) & 0xFFFFFFFF) == 0
    rng = np.random.default_rng(7)
    pf = rng.integers(0, 8, n).astype(np.uint8)
    hc = (rng.random(nq) < 0.5).astype(np.uint8)
    d_pf, d_hc = torch.from_numpy(pf).to(dev), torch.from_numpy(hc).to(dev)
    d_qoff = torch.empty(nq + 1, dtype=torch.int32, device=dev)
    d_qmsg = torch.empty(total, dtype=torch.int32, device=dev)
    d_st = torch.zeros(1, dtype=torch.int32, device=dev)
    d_cnt = torch.empty(n, dtype=torch.int32, device=dev)
    d_fl = torch.empty(n, dtype=torch.uint8, device=dev)

    def step():
        gf.fanout_device(d_off.data_ptr(), d_ids.data_ptr(), n, nq, d_qoff.data_ptr(), d_qmsg.data_ptr(), total, d_st.data_ptr(), st)
        gf.publish_flags_device(d_off.data_ptr(), d_ids.data_ptr(), d_dur.data_ptr(), d_pf.data_ptr(), n, True, d_hc.data_ptr(), nq,
                                d_cnt.data_ptr(), d_fl.data_ptr(), st)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    l0 = L.chq_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    launches = L.chq_kernel_launches() - l0
    assert int(d_st[0]) == 0
    q_off = d_qoff.cpu().numpy().view(np.uint32)
    lens = np.diff(q_off.astype(np.int64))
    # parity + CPU baseline on the first --cpu-sample messages (their own transpose: a prefix of every queue's list)
    ns = min(a.cpu_sample, n)
    h_off = d_off[: ns + 1].cpu().numpy().view(np.uint32)
    h_ids = d_ids[: int(h_off[-1])].cpu().numpy().view(np.uint32)
    t0 = time.perf_counter()
    w_off, w_msg = po.fanout_transpose(h_off, h_ids, nq)
    cpu_s = time.perf_counter() - t0
    q_msg = d_qmsg.cpu().numpy().view(np.uint32)
    chk = rng.integers(0, nq, 2000)
    for qq in chk:   # the sample's list is the head of the full list (publish order)
        want = w_msg[w_off[qq]:w_off[qq + 1]]
        got = q_msg[q_off[qq]:q_off[qq + 1]]
        assert np.array_equal(got[: len(want)], want) and np.all(np.diff(got.astype(np.int64)) > 0)
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    alg = 2 * 4 * total + 2 * 4 * total + 3 * 4 * nq + n * (4 + 4 + 1 + 1 + 1)
    print(json.dumps({
        "metric": "messages/sec through the fan-out transpose (queue -> messages in publish order) + publish flags",
        "value": n / (ms * 1e-3), "unit": "msgs/s", "ms_per_step": ms, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
        "higher_is_better": True, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "routed batch of %d keys against %d bindings (T(seed 4), COMPAT): %d ids into %d queues; "
                               "longest queue list %d, queues with > 1024 messages %d, empty queues %d"
                               % (n, a.bindings, total, nq, int(lens.max()), int((lens > 1024).sum()), int((lens == 0).sum())),
                   "ids_per_second": total / (ms * 1e-3)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_step": alg,
                     "note": "one L2 atomic per id in the count and one in the scatter: atomic throughput, not HBM, sets the pace"},
        "cpu_baseline": {"value": ns / cpu_s, "unit": "msgs/s", "cores": 1, "kind": "port",
                         "sample": "numpy stable argsort over the %d ids of the first %d messages, %.2f s (the reference appends to "
                                   "immutable Vectors inside a Map, ExchangeEntity.scala:305-307; no JVM here)" % (len(h_ids), ns, cpu_s)},
    }))


if __name__ == "__main__":
    main()
