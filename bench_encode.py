#!/usr/bin/env python
"""Secondary benchmark: delivery-side encode — one Basic.Deliver per message of the config-5
stream (1 GiB of publishes with 4 KB bodies), rendered on the device from the decoded batch.

  python bench_encode.py [--bytes N] [--segments C] [--steps K] [--warmup W] [--frame-max F]

Prints one JSON line: deliveries/s and output GB/s of chq_deliver_encode_device (CUDA events; input batch and
records resident in HBM), the HBM roofline fraction (bodies read once, stream written once) and the restated
AMQCommand.render in plain Python on a small sample (the parity check; an interpreter is not a fair CPU figure)."""
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
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--segments", type=int, default=1024)
    ap.add_argument("--body", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frame-max", type=int, default=131072)
    ap.add_argument("--cpu-sample", type=int, default=3000)
    a = ap.parse_args()
    import torch

    from chanamq_b200 import decode as gd
    from chanamq_b200 import encode as ge
    from chanamq_b200._lib import load_library
    from chanamq_b200.workload import TopicWorkload, amqp_stream
    from oracle import pyoracle as po

    if not torch.cuda.is_available():
        raise SystemExit("bench_encode.py: no CUDA device; the product has no CPU path")
    L = load_library()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    data, seg, n_msgs = amqp_stream(TopicWorkload(5), a.bytes, a.segments, body_bytes=a.body, stream=5)
    d_bytes = torch.from_numpy(data).to(dev)
    d_seg = torch.from_numpy(seg.view(np.int64).copy()).to(dev)
    cap = n_msgs + n_msgs // 500 + 1024
    d_recs = torch.zeros(cap * 8, dtype=torch.int64, device=dev)
    d_misc = torch.zeros(4, dtype=torch.int64, device=dev)
    d_cons = torch.zeros(a.segments, dtype=torch.int64, device=dev)
    d_err = torch.zeros(a.segments, dtype=torch.int32, device=dev)
    d_eoff = torch.zeros(a.segments, dtype=torch.int64, device=dev)
    gd.decode_device(d_bytes.data_ptr(), len(data), d_seg.data_ptr(), a.segments, d_recs.data_ptr(), cap, d_misc.data_ptr(),
                     d_cons.data_ptr(), d_err.data_ptr(), d_eoff.data_ptr(), d_misc.data_ptr() + 8, st)
    torch.cuda.synchronize()
    n = int(d_misc[0])
    assert n == n_msgs and (int(d_misc[1]) & 0xFFFFFFFF) == 0
    # one delivery per message: consumer tags "amq.ctag-<k>", delivery tags 1..n, channels 1..64
    tags = [("amq.ctag-%03d" % k).encode() for k in range(64)]
    tag_off = np.cumsum([0] + [len(t) for t in tags])
    k = np.arange(n) % 64
    darr = np.zeros(n, dtype=ge.DELIVERY_DTYPE)
    darr["rec"], darr["delivery_tag"] = np.arange(n), np.arange(1, n + 1)
    darr["ctag_off"], darr["ctag_len"], darr["channel"] = tag_off[k], len(tags[0]), 1 + k
    d_tags = torch.from_numpy(np.frombuffer(b"".join(tags), dtype=np.uint8).copy()).to(dev)
    d_del = torch.from_numpy(np.frombuffer(darr.tobytes(), dtype=np.uint8).copy()).to(dev)
    out_cap = len(data) + 128 * n
    d_out = torch.empty(out_cap, dtype=torch.uint8, device=dev)
    d_off = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    d_st = torch.zeros(1, dtype=torch.int32, device=dev)

    def step():
        ge.deliver_encode_device(d_bytes.data_ptr(), d_recs.data_ptr(), n, d_tags.data_ptr(), d_del.data_ptr(), n, a.frame_max,
                                 d_out.data_ptr(), out_cap, d_off.data_ptr(), d_st.data_ptr(), st)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    assert int(d_st[0]) == 0
    l0 = L.chq_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    launches = L.chq_kernel_launches() - l0
    report(a, po, gd, data, d_recs, d_off, d_out, darr, tags, k, n, ms, launches)


def report(a, po, gd, data, d_recs, d_off, d_out, darr, tags, k, n, ms, launches):
    """parity of a sample against the restated render, then the JSON line"""
    recs = np.frombuffer(d_recs[: n * 8].cpu().numpy().tobytes(), dtype=gd.REC_DTYPE)
    off = d_off.cpu().numpy()
    out_bytes = int(off[-1])
    ns = min(a.cpu_sample, n)
    head = d_out[: int(off[ns])].cpu().numpy().tobytes()
    raw = data[: int(recs["body_off"][ns - 1] + recs["body_size"][ns - 1]) + 8].tobytes() if ns else b""
    t0 = time.perf_counter()
    for i in range(ns):
        r = recs[i]
        p = int(r["method_off"])
        p += 8 + int.from_bytes(raw[p + 3:p + 7], "big")                  # the HEADER frame follows the METHOD frame
        hl = int.from_bytes(raw[p + 3:p + 7], "big")
        body = raw[int(r["body_off"]):int(r["body_off"]) + int(r["body_size"])]   # the stream has single-frame bodies
        want = po.render_deliver(int(darr["channel"][i]), tags[k[i]], int(darr["delivery_tag"][i]), False,
                                 raw[int(r["exch_off"]):int(r["exch_off"]) + int(r["exch_len"])],
                                 raw[int(r["key_off"]):int(r["key_off"]) + int(r["key_len"])], raw[p + 7:p + 7 + hl], body, a.frame_max)
        assert head[off[i]:off[i + 1]] == want, i
    cpu_s = time.perf_counter() - t0
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    alg = int(recs["body_size"].sum()) + out_bytes + n * (64 + 24 + 8 + 60)   # bodies in, stream out, records / deliveries / offsets / headers
    print(json.dumps({
        "metric": "deliveries/sec rendered (Basic.Deliver + content header + body frames, %d-byte bodies)" % a.body,
        "value": n / (ms * 1e-3), "unit": "deliveries/s", "ms_per_step": ms, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
        "higher_is_better": True, "dtype": "u8", "data": "synthetic", "out_gb_per_s": out_bytes / (ms * 1e-3) / 1e9,
        "config": {"workload": "%d deliveries, one per message of the config-5 stream (%d bytes in %d segments), frame_max %d, %d "
                               "output bytes" % (n, len(data), a.segments, a.frame_max, out_bytes)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / peak,
                     "algorithmic_bytes_per_step": alg},
        "cpu_baseline": {"value": ns / cpu_s if ns else None, "unit": "deliveries/s", "cores": 1, "kind": "port",
                         "sample": "%d deliveries through the restated AMQCommand.render in plain Python (the parity check; an "
                                   "interpreter, not a fair CPU figure), %.2f s" % (ns, cpu_s)},
    }))


if __name__ == "__main__":
    main()
