#!/usr/bin/env python
"""Secondary benchmark: the decode half of the hot path (BASELINE.json configs[4]: raw AMQP frame stream,
4 KB bodies, 1 GiB byte batch), optionally followed by routing on the device.

  python bench_decode.py [--bytes N] [--segments C] [--steps K] [--warmup W] [--route]

Prints one JSON line: messages/s and stream GB/s of chq_decode_device with the bytes resident in HBM
(CUDA events), the algorithmic bytes (8 per frame + method + header payloads + one 64-byte
record per message — bodies are not algorithmic), and the CPU restatement's rate on a sample.
bench.py stays the headline (routing); this line documents the decode stage."""
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
    ap.add_argument("--cpu-sample-bytes", type=int, default=256 << 20)
    ap.add_argument("--route", action="store_true",
                    help="config 5 as written: decode, then route the decoded keys on the device against config 2's table "
                         "(100 k bindings, mixed */# patterns, SPEC semantics); also times the chain with the stream coming from host memory")
    a = ap.parse_args()
    import torch

    from chanamq_b200 import decode as gd
    from chanamq_b200._lib import load_library
    from chanamq_b200.workload import TopicWorkload, amqp_stream
    from oracle import pyoracle as po

    if not torch.cuda.is_available():
        raise SystemExit("bench_decode.py: no CUDA device; the product has no CPU path")
    L = load_library()
    dev = torch.device("cuda:0")
    w = TopicWorkload(5)
    t0 = time.perf_counter()
    data, seg, n_msgs = amqp_stream(w, a.bytes, a.segments, body_bytes=a.body, stream=5)
    gen_s = time.perf_counter() - t0
    n_bytes = len(data)
    d_bytes = torch.from_numpy(data).to(dev)
    d_seg = torch.from_numpy(seg.view(np.int64).copy()).to(dev)
    cap = n_msgs + n_msgs // 500 + 1024
    d_recs = torch.zeros(cap * 8, dtype=torch.int64, device=dev)
    d_misc = torch.zeros(4, dtype=torch.int64, device=dev)
    d_cons = torch.zeros(a.segments, dtype=torch.int64, device=dev)
    d_err = torch.zeros(a.segments, dtype=torch.int32, device=dev)
    d_eoff = torch.zeros(a.segments, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def step():
        gd.decode_device(d_bytes.data_ptr(), n_bytes, d_seg.data_ptr(), a.segments, d_recs.data_ptr(), cap, d_misc.data_ptr(),
                         d_cons.data_ptr(), d_err.data_ptr(), d_eoff.data_ptr(), d_misc.data_ptr() + 8, st)

    route = None
    if a.route:
        from chanamq_b200 import CHQ_MODE_SPEC, CHQ_TOPIC, GpuMatcher
        from chanamq_b200.workload import queue_names_packed

        w2 = TopicWorkload(5)   # the stream's keys come from T(seed 5): bind patterns of the same distribution
        pb, poff, q, d = w2.patterns(100_000, hash_tail=True)
        qb, qoff = queue_names_packed(q)
        m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
        m.bind_many(pb, poff, qb, qoff, d)
        m.commit()
        ids_cap = 64 * cap
        d_koff = torch.zeros(cap, dtype=torch.int32, device=dev)
        d_klen = torch.zeros(cap, dtype=torch.int32, device=dev)
        d_ridx = torch.zeros(cap, dtype=torch.int32, device=dev)
        d_offsets = torch.zeros(cap + 1, dtype=torch.int32, device=dev)
        d_ids = torch.zeros(ids_cap, dtype=torch.int32, device=dev)
        d_dur = torch.zeros(cap, dtype=torch.uint8, device=dev)
        d_kerr = torch.zeros(cap, dtype=torch.uint8, device=dev)
        d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
        decode_only = step

        def step():   # noqa: F811 — decode, compact the keys, route them; n_msgs is known to the bench, a broker reads d_misc[2]
            decode_only()
            gd.publish_keys_device(d_recs.data_ptr(), d_misc.data_ptr(), cap, d_koff.data_ptr(), d_klen.data_ptr(), d_ridx.data_ptr(),
                                   d_misc.data_ptr() + 16, st)
            m.route_keys_device(d_bytes.data_ptr(), d_koff.data_ptr(), d_klen.data_ptr(), n_msgs, d_offsets.data_ptr(), d_ids.data_ptr(),
                                ids_cap, d_dur.data_ptr(), d_kerr.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)

        route = True
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
    e2e = None
    if route:
        assert (int(d_tot[1]) & 0xFFFFFFFF) == 0 and (int(d_misc[2]) & 0xFFFFFFFF) == n_msgs
        routed_ids = int(d_tot[0])
        # end to end: the stream starts in pinned host memory, the CSR (offsets + ids) ends in host memory
        h_bytes = torch.from_numpy(data).pin_memory()
        h_off = torch.empty(n_msgs + 1, dtype=torch.int32).pin_memory()
        h_ids = torch.empty(routed_ids, dtype=torch.int32).pin_memory()
        k_e2e = max(2, min(a.steps, 5))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            d_bytes.copy_(h_bytes, non_blocking=True)
            step()
            h_off.copy_(d_offsets[: n_msgs + 1], non_blocking=True)
            h_ids.copy_(d_ids[:routed_ids], non_blocking=True)
            torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / k_e2e
        e2e = {"value": n_msgs / (e2e_ms * 1e-3), "unit": "msgs/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(n_bytes),
               "d2h_bytes_per_step": int(4 * (n_msgs + 1) + 4 * routed_ids), "routed_ids_per_step": routed_ids,
               "timing": "host wall clock: 1 GiB stream H2D from pinned memory, decode + key compaction + routing, CSR D2H"}
    n_recs, status = int(d_misc[0]), int(d_misc[1]) & 0xFFFFFFFF
    assert status == 0 and n_recs == n_msgs and int(d_err.abs().sum()) == 0, (status, n_recs, n_msgs)
    recs = np.frombuffer(d_recs[: n_recs * 8].cpu().numpy().tobytes(), dtype=gd.REC_DTYPE)
    # algorithmic bytes: frame headers + method and header payloads + records
    hdr_payload = 14 + 1 + np.where(recs["flags"] & 16, recs["hdr_tbl_len"].astype(np.int64) + 4 + 8, 0)
    method_payload = 6 + 1 + recs["exch_len"].astype(np.int64) + 1 + recs["key_len"].astype(np.int64) + 1
    alg = int(3 * 8 * n_recs + hdr_payload.sum() + method_payload.sum() + 64 * n_recs)
    # CPU restatement on a sample of whole segments
    n_seg_s = max(1, int(a.segments * min(1.0, a.cpu_sample_bytes / n_bytes)))
    sb = data[: int(seg[n_seg_s])]
    t0 = time.perf_counter()
    c_recs, _, c_err, _ = po.decode(sb, seg[: n_seg_s + 1])
    cpu_s = time.perf_counter() - t0
    assert c_recs.tobytes() == recs[: len(c_recs)].tobytes() and not c_err.any()
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    print(json.dumps({
        "metric": ("decoded + routed msgs/sec (raw AMQP frame stream, 4 KB bodies, 100 k bindings)" if route
                   else "decoded msgs/sec (raw AMQP frame stream, 4 KB bodies)"), "value": n_msgs / (ms * 1e-3), "unit": "msgs/s",
        "e2e": e2e,
        "ms_per_step": ms, "stream_gb_per_s": n_bytes / (ms * 1e-3) / 1e9, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
        "config": {"workload": "AMQP publish stream, %d bytes in %d connection segments, %d-byte bodies, %d messages (T(seed 5) keys, "
                               "10%% with expiration + 8-field headers table, 0.1%% heartbeats)" % (n_bytes, a.segments, a.body, n_msgs),
                   "generation_s": gen_s},
        "gpu_launches": int(launches), "dtype": "u8", "data": "synthetic",
        "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_step": alg,
                     "note": "serial frame walk per connection segment (pointer chase): latency-bound, not bandwidth-bound"},
        "cpu_baseline": {"value": len(c_recs) / cpu_s, "unit": "msgs/s", "cores": 1, "kind": "port",
                         "sample": "%d segments (%d bytes) through the restated FrameParser/CommandAssembler on 1 thread, %.2f s; "
                                   "stream %.2f GB/s" % (n_seg_s, len(sb), cpu_s, len(sb) / cpu_s / 1e9)},
    }))


if __name__ == "__main__":
    main()
