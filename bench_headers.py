#!/usr/bin/env python
"""Secondary benchmark: headers exchange, x-match any/all (BASELINE.json configs[2]: 50 k bindings x 8
header fields, 10 M messages = 10 batches of 2^20).  SPEC mode only: the reference has no headers matcher
(chana-mq-server/.../entity/ExchangeEntity.scala:210), so parity is against the brute-force oracle
(oracle/headers_spec.cpp, "parity unpinned") and the CPU baseline is that same brute force on a small sample.

  python bench_headers.py [--bindings P] [--batch N] [--steps K] [--warmup W]

Prints one JSON line: messages/s of chq_route_device with the headers tables resident in HBM (CUDA events),
the end-to-end rate through chq_route (host tables in, host CSR out), per-kernel times, and the roofline
fraction from SURVEY.md §8(d): B_hdr = N (8 F + 4 + 4 M + 4) + P (8 F + 8), F = 8.
bench.py stays the headline (topic routing); this line documents config 3."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
N_BATCHES = 10


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bindings", type=int, default=50_000)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--cpu-sample", type=int, default=1500)
    a = ap.parse_args()
    import torch

    from chanamq_b200 import CHQ_HEADERS, CHQ_MODE_SPEC, GpuMatcher
    from chanamq_b200._lib import load_library
    from chanamq_b200.workload import HeadersWorkload
    from oracle import pyoracle as po

    if not torch.cuda.is_available():
        raise SystemExit("bench_headers.py: no CUDA device; the product has no CPU path")
    L = load_library()
    dev = torch.device("cuda:0")
    w = HeadersWorkload(3)
    bb, bo, q, d = w.bindings(a.bindings)
    binds = [("q%06d" % q[i], bool(d[i]), bytes(bb[bo[i]:bo[i + 1]])) for i in range(a.bindings)]
    m = GpuMatcher(CHQ_HEADERS, CHQ_MODE_SPEC)
    t0 = time.perf_counter()
    for qn, du, args in binds:
        assert m.bind_args("", qn, du, args) == 0
    m.commit()
    build_s = time.perf_counter() - t0
    n = a.batch
    batches = [w.messages(n, stream=10 + b) for b in range(N_BATCHES)]
    d_blob = [torch.from_numpy(b[0]).to(dev) for b in batches]
    d_off = [torch.from_numpy(b[1].view(np.int32).copy()).to(dev) for b in batches]
    cap = 64 * n
    d_offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_ids = torch.empty(cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    d_err = torch.empty(n, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def step(i):
        b = i % N_BATCHES
        m.route_device(d_blob[b].data_ptr(), d_off[b].data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), cap,
                       d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)

    for i in range(a.warmup):
        step(i)
    torch.cuda.synchronize()
    assert (int(d_tot[1]) & 0xFFFFFFFF) == 0, "route status %d" % int(d_tot[1])
    m.profile_kernels(True)
    l0 = L.chq_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step(a.warmup + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    launches = L.chq_kernel_launches() - l0
    kms, prof_n = m.profile_kernels(False)
    assert (int(d_tot[1]) & 0xFFFFFFFF) == 0
    total_ids = int(d_tot[0])
    mbar = total_ids / n   # of the last batch routed

    # end to end: host tables in, host CSR out, one blocking chq_route per batch
    t0 = time.perf_counter()
    e2e_steps = max(3, min(a.steps, 10))
    chk = 0
    for i in range(e2e_steps):
        r = m.route_headers_packed(batches[i % N_BATCHES][0], batches[i % N_BATCHES][1], copy=False)
        chk += int(r.offsets[-1])
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    h2d = int(np.mean([len(b[0]) + 4 * (n + 1) for b in batches]))
    d2h = int(4 * (n + 1) + 4 * mbar * n + 2 * n)

    # parity on a sample + CPU baseline: the brute-force oracle (the reference has no algorithm to restate)
    o = po.HeadersOracle()
    for qn, du, args in binds:
        o.bind(qn, du, args)
    ns = min(a.cpu_sample, n)
    blob0, off0 = batches[0]
    t0 = time.perf_counter()
    w_off, w_ids, w_dur, _ = o.route_packed(blob0[: off0[ns]], off0[: ns + 1])
    cpu_s = time.perf_counter() - t0
    r0 = m.route_headers_packed(blob0, off0)
    assert np.array_equal(r0.offsets[: ns + 1], w_off) and np.array_equal(r0.queue_ids[: w_off[-1]], w_ids)

    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    F = 8
    alg = n * (8 * F + 4 + 4 * mbar + 4) + a.bindings * (8 * F + 8)
    stream_bytes = float(np.mean([len(b[0]) for b in batches]))
    print(json.dumps({
        "metric": "routed msgs/sec (headers x-match any/all, %d bindings x 8 fields)" % a.bindings,
        "value": n / (ms * 1e-3), "unit": "msgs/s", "ms_per_step": ms, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
        "higher_is_better": True, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "headers exchange, %d bindings x 8 argument slots (half any, half all with 0-6 void arguments), "
                               "messages of 8 fields from 100 k Zipf templates (seed 3), batches of %d x %d resident batches, "
                               "AMQP x-match semantics (SPEC; parity unpinned)" % (a.bindings, n, N_BATCHES),
                   "mean_queues_per_message": mbar, "table_build_s": build_s,
                   "headers_bytes_per_message": stream_bytes / n},
        "e2e": {"value": n / (e2e_ms * 1e-3), "unit": "msgs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms, "checksum_ids": chk,
                "timing": "host wall clock around blocking chq_route calls, host tables in, host CSR out"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_step": alg,
                     "per_kernel_ms": {("k_headers_match" if k == "k_tokenize" else k): v / max(1, prof_n) for k, v in kms.items()},
                     "raw_stream_gb_per_s": stream_bytes / (ms * 1e-3) / 1e9,
                     "note": "the raw tables are %.0f bytes per message, 2x the interned (name, value) pairs the formula counts"
                             % (stream_bytes / n)},
        "cpu_baseline": {"value": ns / cpu_s, "unit": "msgs/s", "cores": 1, "kind": "port",
                         "sample": "%d messages of batch 0 against all %d bindings, brute force over (message, binding) pairs on "
                                   "1 thread, %.1f s; the reference has no headers matcher to restate" % (ns, a.bindings, cpu_s)},
    }))


if __name__ == "__main__":
    main()
