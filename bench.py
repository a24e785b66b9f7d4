#!/usr/bin/env python
"""Benchmark of the routing hot path: routed msgs/sec, topic exchange, 1 M bindings (BASELINE.json).

One "step" = one batch of 2^20 synthetic routing keys routed against the 1 M-binding table
(tokenise + fused route kernel).  Ten distinct batches (10.5 M keys, the "10 M messages" of
BASELINE.json configs[3]) stay resident in HBM and rotate, so no step re-reads what the previous one
left in L2, and every step writes ~0.4 GB of queue ids (> the 126 MB L2).

  python bench.py [--gpus N] [--steps K] [--warmup W]          product arm (CUDA, C ABI)
  python bench.py --impl reference ...                         the reference's CPU routing (C++ restatement
                                                               of TrieMatcher: no JVM in this image), all host threads

Multi-GPU (torchrun, one rank per GPU): the path shards by KEYS — every rank holds the whole compiled
table (tens of MB) and routes its own batches; there is no data-path collective, so scaling is weak
(per-GPU work fixed).  The binding-slice mode of BASELINE.json configs[3] is a separate code path
(chanamq_b200.sharded) and is not what this line times.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNIT = "msgs/s"
N_BATCHES = 10


def seed_of(bindings):
    """config 1 = T(seed 1, P = 1 000), config 2 = T(seed 2, P = 100 000), configs 4 (and the headline) =
    T(seed 4, P = 1 000 000)"""
    return {1_000: 1, 100_000: 2}.get(bindings, 4)


def metric_name(a):
    p = a.bindings
    return "routed msgs/sec (topic, %s bindings)" % ("1M" if p == 1_000_000 else ("%dk" % (p // 1000) if p % 1000 == 0 else str(p)))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bindings", type=int, default=1_000_000)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--mode", default="compat", choices=["compat", "spec"],
                    help="compat = the reference's semantics ('#' is a word); spec = AMQP '#' wildcard")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip the short legs for BASELINE.json configs 2, 3 and 5 that the default 1-GPU run appends")
    ap.add_argument("--no-numa", action="store_true", help="leave the process where the launcher put it (default: move it to its GPU's NUMA node)")
    ap.add_argument("--cpu-sample", type=int, default=100_000)
    ap.add_argument("--ref-sample", type=int, default=0, help="keys per reference-arm step (0 = sized from a pilot so that the run takes about 75 s)")
    return ap.parse_args()


def workload_name(a):
    return ("topic exchange, %d bindings with mixed */# patterns (T(seed=%d), one queue per binding), %s semantics, "
            "batches of %d keys x %d resident batches" % (
                a.bindings, seed_of(a.bindings), "reference (COMPAT: '#' literal)" if a.mode == "compat" else "AMQP (SPEC: '#' wildcard)",
                a.batch, N_BATCHES))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_table_inputs(a):
    from chanamq_b200.workload import TopicWorkload, queue_names_packed

    w = TopicWorkload(seed_of(a.bindings))
    pb, poff, q, d = w.patterns(a.bindings, a.bindings, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    return w, (pb, poff, qb, qoff, d)


# --------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the C++ restatement of the reference's TrieMatcher (oracle/)
# --------------------------------------------------------------------------------------------
def cpu_route_rate(a, w, binds, n_keys, threads, steps, warmup, stream0=1000, budget_s=None):
    """n_keys per step; with budget_s the per-step sample is sized from a pilot so the whole run takes ~budget_s"""
    from oracle import pyoracle as po

    o = po.Oracle(po.TOPIC, po.COMPAT if a.mode == "compat" else po.SPEC, inplace=True)
    t0 = time.perf_counter()
    assert o.bind_many(*binds) == 0
    build_s = time.perf_counter() - t0
    if budget_s:
        blob, off = w.keys(2000 * threads, stream=stream0 - 1)
        t0 = time.perf_counter()
        o.route_packed(blob, off, threads)
        pilot_rate = 2000 * threads / (time.perf_counter() - t0)
        n_keys = int(min(1 << 20, max(2000, pilot_rate * budget_s / (steps + warmup))))
    secs, ids, last = [], 0, None
    for s in range(warmup + steps):
        blob, off = w.keys(n_keys, stream=stream0 + s)
        t0 = time.perf_counter()
        o_off, o_ids, o_dur, o_err = o.route_packed(blob, off, threads)
        dt = time.perf_counter() - t0   # lookups + collapsing to the CSR the exchange consumes
        last = (o_off, o_ids, o_dur, o_err)
        if s >= warmup:
            secs.append(dt)
            ids += len(o_ids)
    return n_keys * steps / sum(secs), sum(secs) / steps, build_s, ids / max(1, n_keys * steps), n_keys, last


def run_reference(a, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    w, binds = make_table_inputs(a)
    rate, per_step, build_s, mbar, n_keys, _ = cpu_route_rate(a, w, binds, a.ref_sample, threads, a.steps, a.warmup,
                                                              budget_s=None if a.ref_sample else 75.0)
    sample = "%d keys per step x %d steps on %d threads (keys sharded over a read-only trie); trie build %.1f s not timed" % (
        n_keys, a.steps, threads, build_s)
    print(json.dumps({
        "impl": "reference", "metric": metric_name(a), "value": rate, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(a), "mean_queues_per_message": mbar},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "C++ restatement of the reference's TrieMatcher + ExchangeEntity dedup (oracle/); the Scala itself "
                "cannot run here (no JVM) and is expected to be slower (GC, boxing)",
    }))


# --------------------------------------------------------------------------------------------
# product arm
# --------------------------------------------------------------------------------------------
def run_ours(a, rank, world, local_rank):
    import torch

    from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC, GpuMatcher
    from chanamq_b200._lib import CHQ_IDS_U24, CHQ_IDS_U32, load_library

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)

    L = load_library()
    # this process on its GPU's NUMA node before anything is allocated (pinned staging and result buffers follow it)
    numa_node, numa_cpus = C.c_int32(-1), C.c_int32(0)
    if not a.no_numa:
        L.chq_numa_bind_to_device(local_rank, C.byref(numa_node), C.byref(numa_cpus))
    w, binds = make_table_inputs(a)
    t0 = time.perf_counter()
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT if a.mode == "compat" else CHQ_MODE_SPEC, device=local_rank)
    m.bind_many(*binds)
    m.commit()
    build_s = time.perf_counter() - t0
    if os.environ.get("CHQ_MAX_SMALL"):  # experiment switch: where the staged merge hands over to warp_union_long
        m.debug_limits(0, int(os.environ["CHQ_MAX_SMALL"]))
    st = m.stats()

    # ---- inputs: N_BATCHES distinct batches per rank, in pinned host memory and resident in HBM ----
    n = a.batch
    host, devb = [], []
    words_total = 0
    for b in range(N_BATCHES):
        blob, off = w.keys(n, stream=rank * 1000 + b)
        hb = torch.empty(len(blob), dtype=torch.uint8).pin_memory()
        ho = torch.empty(n + 1, dtype=torch.int32).pin_memory()
        hb.numpy()[:] = blob
        ho.numpy()[:] = off.view(np.int32)
        host.append((hb, ho))
        devb.append((hb.to(dev), ho.to(dev)))
        words_total += int(np.count_nonzero(blob == ord("."))) + n
    wbar = words_total / (N_BATCHES * n)

    # ---- device-resident arm -------------------------------------------------------------------
    ids_cap = int(n * 160)
    d_offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_ids = torch.empty(ids_cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    d_err = torch.empty(n, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2 * N_BATCHES, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device(i):
        b = i % N_BATCHES
        db, do = devb[b]
        m.route_device(db.data_ptr(), do.data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), ids_cap,
                       d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr() + 16 * b, d_tot.data_ptr() + 16 * b + 8, stream)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(a.warmup):
        step_device(i)
    barrier()
    m.profile(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = L.chq_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step_device(a.warmup + i)
    e1.record()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    launches = L.chq_kernel_launches() - launches0
    kms, prof_n = m.profile_kernels(False)
    write_ms = kms["k_write"] + kms["k_write_rest"]   # the pair that writes the ids: the common kinds of message, then the rest
    tok_ms, route_ms = kms["k_tokenize"], kms["k_probe"] + kms["k_offsets"] + write_ms
    tot = d_tot.cpu().numpy().reshape(N_BATCHES, 2)
    used = sorted({(a.warmup + i) % N_BATCHES for i in range(a.steps)} | {i % N_BATCHES for i in range(a.warmup)})
    if any((int(tot[b, 1]) & 0xFFFFFFFF) != 0 for b in used):
        raise SystemExit("bench.py: route status %s" % [int(tot[b, 1]) & 0xFFFFFFFF for b in used])
    ids_per_batch = float(np.mean([tot[b, 0] for b in used]))
    mbar = ids_per_batch / n

    # ---- end to end through the host-buffer C-ABI call (pinned inputs, H2D + kernels + D2H) -----
    depth = 3

    def e2e_run(steps, base):
        inflight = []
        chk = 0
        for i in range(steps):
            slot = i % depth
            if len(inflight) == depth:
                r = m.wait(inflight.pop(0))
                chk += int(r.offsets[-1])
            hb, ho = host[(base + i) % N_BATCHES]
            m.submit(slot, hb.data_ptr(), ho.data_ptr(), n)
            inflight.append(slot)
        for s in inflight:
            r = m.wait(s)
            chk += int(r.offsets[-1])
        return chk

    e2e_run(min(a.warmup, depth) + depth, 0)
    barrier()
    t0 = time.perf_counter()
    chk = e2e_run(a.steps, a.warmup)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    # the same calls with the ids delivered as 3-byte values (chq_set_ids_format: optional, tables below 2^24 queues)
    m.set_ids_format(CHQ_IDS_U24)
    e2e_run(depth, 0)
    barrier()
    t0 = time.perf_counter()
    chk24 = e2e_run(a.steps, a.warmup)
    torch.cuda.synchronize()
    e2e24_s = time.perf_counter() - t0
    m.set_ids_format(CHQ_IDS_U32)
    if chk24 != chk:
        raise SystemExit("bench.py: the packed result counts %d ids, the uint32 one %d" % (chk24, chk))
    clocks = sampler.stop()  # sampled from the start of the device-timed region to the end of the end-to-end one
    clocks["window"] = "device-resident + end-to-end timed regions"
    h2d = int(np.mean([hb.numel() + ho.numel() * 4 for hb, ho in host]))
    d2h = int((n + 1) * 4 + 2 * n + 16 + ids_per_batch * 4)

    slices = run_binding_slices(a, rank, world, local_rank, w, binds, mbar, dist, barrier) if world > 1 else None

    t = torch.tensor([dev_ms, e2e_s * 1e3, e2e24_s * 1e3], dtype=torch.float64, device=dev)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, e2e24_ms = float(t[0]), float(t[1]), float(t[2])
    value = world * n * a.steps / (dev_ms * 1e-3)
    e2e_value = world * n * a.steps / (e2e_ms * 1e-3)

    if rank == 0:
        peaks, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        try:
            peaks = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            peak_src = "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
        wp = st["pattern_words"] / max(1, st["bindings"])
        # B_topic = N(4 W + 4 + 4 M + 4) + P(4 Wp + 8), per batch, charged to the three route kernels
        # (k_probe + k_offsets + k_write: one launch group per batch); k_write alone moves N(4 M + 4 + 1) + 4 P of it
        alg_bytes = n * (4 * wbar + 4 + 4 * mbar + 4) + st["bindings"] * (4 * wp + 8)
        route_ms_avg = route_ms / max(1, prof_n)
        write_bytes = n * (4 * mbar + 4 + 1) + 4 * st["bindings"]
        achieved = alg_bytes / (route_ms_avg * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "route_traffic.json")))["dram_bytes_per_launch"]
        except Exception:
            pass
        out = {
            "metric": metric_name(a), "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32", "data": "synthetic",
            "config": {"workload": workload_name(a), "mean_words_per_key": wbar, "mean_queues_per_message": mbar,
                       "sharding": "keys (table replicated per GPU, no data-path collective)",
                       "l2": "10 distinct resident batches rotate: %.0f MB of keys in all and ~%.0f MB of ids written per step, so "
                             "a batch's keys have left the 126 MB L2 when its turn comes again (inputs larger than L2, no flush)" % (
                                 sum(hb.numel() + 4 * ho.numel() for hb, ho in host) / 1e6, ids_per_batch * 4 / 1e6),
                       "table": {k: st[k] for k in ("bindings", "queues", "groups", "shapes", "words", "wstride", "table_bytes")},
                       "table_build_s": build_s,
                       "numa": {"node_rank0": numa_node.value, "cpus_rank0": numa_cpus.value} if not a.no_numa else None},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / a.steps, "pipeline_depth": depth, "checksum_ids": chk,
                    "timing": "host wall clock around K pipelined chq_route_submit/chq_route_wait calls, pinned host keys in, host CSR out"},
            "e2e_ids_u24": {"value": world * n * a.steps / (e2e24_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e24_ms / a.steps,
                            "d2h_bytes_per_step": int((n + 1) * 4 + 2 * n + 16 + ids_per_batch * 3),
                            "what": "the e2e calls with chq_set_ids_format(CHQ_IDS_U24): queue ids reach the host as 3-byte "
                                    "values (one more kernel, k_pack24, per batch); optional, `e2e` above is the uint32 form"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks, "unit": "GB/s", "frac": achieved / peaks,
                         "traffic": traffic, "kernel": "route group = k_probe + k_offsets + k_write + k_write_rest (one each per batch)",
                         "kernel_ms": route_ms_avg,
                         "per_kernel_ms": {k: v / max(1, prof_n) for k, v in kms.items()},
                         "dominant": {"kernel": "k_write + k_write_rest", "ms": write_ms / max(1, prof_n),
                                      "algorithmic_bytes_per_launch": write_bytes,
                                      "achieved": write_bytes / (write_ms / max(1, prof_n) * 1e-3) / 1e9,
                                      "frac": write_bytes / (write_ms / max(1, prof_n) * 1e-3) / 1e9 / peaks},
                         "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                         "tokenize_ms": tok_ms / max(1, prof_n), "kernel_share_of_step": route_ms / dev_ms if world == 1 else None},
        }
        if slices is not None:
            out["binding_slices"] = slices
        if world == 1 and not a.no_configs:
            out["configs"] = run_config_legs()
        if world == 1 and not a.no_cpu_baseline:
            rate, per_step, cbuild, cm, _, want = cpu_route_rate(a, w, binds, a.cpu_sample, 1, 1, 0, stream0=0)
            # parity of the timed path on the timed configuration: the oracle has just routed the first cpu_sample keys
            # of batch 0; route batch 0 once more through the product and compare that prefix bit for bit
            ns = min(a.cpu_sample, n)
            step_device(0)
            torch.cuda.synchronize()
            g_off = d_offsets[: ns + 1].cpu().numpy().view(np.uint32)
            g_ids = d_ids[: int(g_off[-1])].cpu().numpy().view(np.uint32)
            equal = bool(np.array_equal(g_off, want[0]) and np.array_equal(g_ids, want[1]) and
                         np.array_equal(d_dur[:ns].cpu().numpy(), want[2]) and
                         np.array_equal(d_err[:ns].cpu().numpy() != 0, want[3] != 0))
            out["parity"] = {"checked_keys": int(ns), "checked_ids": int(len(want[1])), "equal": equal,
                             "against": "CPU oracle (restated TrieMatcher + ExchangeEntity dedup), first keys of batch 0"}
            out["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port",
                                   "sample": "%d keys of batch 0 on 1 thread (one exchange = one actor = one thread in the "
                                             "reference, ExchangeEntity.scala:107), %.1f s; trie build %.1f s not timed; "
                                             "C++ restatement of TrieMatcher (no JVM here)" % (a.cpu_sample, per_step, cbuild)}
        print(json.dumps(out))
        if (out.get("binding_slices") or {}).get("parity", {}).get("equal") is False:
            raise SystemExit("bench.py: the binding-slice result differs from the oracle's")
        if out.get("parity", {}).get("equal") is False:
            raise SystemExit("bench.py: the product's result differs from the oracle's on the benchmarked configuration")
    if dist:
        dist.barrier()
        dist.destroy_process_group()


def run_config_legs():
    """Short legs for the other BASELINE.json configs, each in its own process after the headline (the GPU is idle by
    then): config 2 (100 k bindings, SPEC '#' semantics), config 3 (headers x-match), config 5 (decode -> route).  Every
    leg checks a sample of its result against its oracle and exits non-zero on a mismatch (parity = it exited 0)."""
    legs = {
        "config2_topic_100k_spec": [os.path.join(ROOT, "bench.py"), "--bindings", "100000", "--mode", "spec", "--steps", "30",
                                    "--warmup", "3", "--cpu-sample", "20000", "--no-configs"],
        "config3_headers_50k": [os.path.join(ROOT, "bench_headers.py"), "--steps", "10", "--warmup", "3", "--cpu-sample", "300"],
        "config5_decode_route_1gib": [os.path.join(ROOT, "bench_decode.py"), "--route", "--steps", "5", "--warmup", "3",
                                      "--cpu-sample-bytes", str(64 << 20)],
    }
    out = {}
    for name, cmd in legs.items():
        t0 = time.perf_counter()
        try:
            r = subprocess.run([sys.executable] + cmd, capture_output=True, text=True, timeout=240)
            lines = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
            if r.returncode != 0 or not lines:
                out[name] = {"failed": (r.stderr.strip().splitlines() or ["exit %d" % r.returncode])[-1][:300]}
                continue
            d = json.loads(lines[-1])
            par = d.get("parity", {})
            out[name] = {"metric": d.get("metric"), "value": d.get("value"), "unit": d.get("unit"), "ms_per_step": d.get("ms_per_step"),
                         "roofline_frac": (d.get("roofline") or {}).get("frac"), "e2e": (d.get("e2e") or {}).get("value"),
                         "cpu_baseline": (d.get("cpu_baseline") or {}).get("value"),
                         "parity": {"equal": par.get("equal", True), "checked": par.get("checked_keys", "sample asserted inside the leg")},
                         "workload": (d.get("config") or {}).get("workload"), "leg_s": time.perf_counter() - t0}
        except Exception as e:  # a leg must never take the headline line down with it
            out[name] = {"failed": repr(e)[:300]}
    return out


def run_binding_slices(a, rank, world, local_rank, w, binds, mbar, dist, barrier):
    """BASELINE.json configs[3] as written: binding i on GPU i mod G, every GPU routes ALL keys of a batch against its
    slice, the slices' results meet over NVLink (NCCL all-gather of the small arrays + peer reads of the ids) and rank g
    ends with the finished result of messages [g n / G, (g+1) n / G) — one chq_shard_route_device call per rank per batch.
    A step = one batch of n keys for the WHOLE job (the same keys on every rank)."""
    import torch

    from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC
    from chanamq_b200.sharded import ShardedExchange

    dev = torch.device("cuda", local_rank)
    n = a.batch
    box = [ShardedExchange.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    t0 = time.perf_counter()
    ex = ShardedExchange(box[0], rank, world, local_rank, CHQ_TOPIC, CHQ_MODE_COMPAT if a.mode == "compat" else CHQ_MODE_SPEC)
    ex.bind_many(*binds)
    ex.commit()
    mb = torch.tensor([mbar], dtype=torch.float64, device=dev)   # (every rank measured it on its own key batches)
    dist.all_reduce(mb, op=dist.ReduceOp.MAX)
    mbar = float(mb[0])
    slice_cap = int(n * (mbar / world * 1.3 + 8)) + 4096
    pull_cap = int(n / world * (mbar / world * 1.5 + 8)) + 4096
    ids_cap = int(n / world * (mbar * 1.5 + 16)) + 4096
    ex.reserve(n, slice_cap, pull_cap)
    build_s = time.perf_counter() - t0
    keys = []
    for b in range(N_BATCHES):  # the same batches on every rank
        blob, off = w.keys(n, stream=5000 + b)
        keys.append((torch.from_numpy(blob.copy()).to(dev), torch.from_numpy(off.view(np.int32).copy()).to(dev), blob if b == 0 else None,
                     off if b == 0 else None))
    nn_max = n // world + 2
    d_offsets = torch.empty(nn_max + 1, dtype=torch.int32, device=dev)
    d_ids = torch.empty(ids_cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(nn_max, dtype=torch.uint8, device=dev)
    d_err = torch.empty(nn_max, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def step(i):
        db, do, _, _ = keys[i % N_BATCHES]
        return ex.route_device(db.data_ptr(), do.data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), ids_cap, d_dur.data_ptr(),
                               d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, stream)

    for i in range(a.warmup):
        step(i)
    barrier()
    ex.stage_ms(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step(a.warmup + i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    stages, stage_n = ex.stage_ms(False)
    status = int(d_tot[1].item()) & 0xFFFFFFFF
    # parity: batch 0 once more; rank 0 compares the head of its range with the oracle over the WHOLE table
    m0, m1 = step(0)
    torch.cuda.synchronize()
    parity = None
    if rank == 0:
        from oracle import pyoracle as po

        o = po.Oracle(po.TOPIC, po.COMPAT if a.mode == "compat" else po.SPEC, inplace=True)
        assert o.bind_many(*binds) == 0
        ns = min(20_000, m1 - m0)
        blob, off = keys[0][2], keys[0][3]
        w_off, w_ids, w_dur, _ = o.route_packed(blob[: off[ns]], off[: ns + 1], os.cpu_count() or 1)
        g_off = d_offsets[: ns + 1].cpu().numpy().view(np.uint32)
        equal = bool(np.array_equal(g_off, w_off) and np.array_equal(d_ids[: int(w_off[-1])].cpu().numpy().view(np.uint32), w_ids) and
                     np.array_equal(d_dur[:ns].cpu().numpy(), w_dur))
        parity = {"checked_keys": int(ns), "checked_ids": int(len(w_ids)), "equal": equal,
                  "against": "CPU oracle over the whole table, head of rank 0's message range of batch 0"}
    t = torch.tensor([ms, float(status != 0)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    ex.close()
    return {"value": n * a.steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / a.steps, "status_ok": float(t[1]) == 0.0,
            "what": "bindings sliced i mod G; every GPU routes all %d keys of a step against its slice; every owner receives its "
                    "message range of every slice's offsets and durable flags over NCCL (CHQ_SHARD_ALLGATHER=1: the whole "
                    "arrays, all-gathered); each GPU pulls the ids of its own n/G messages from the peers' buffers over NVLink "
                    "and merges them (chq_shard_route_device)" % n,
            "nccl_bytes_received_per_rank_per_step": int((world - 1) * ((n // world + 1) * 4 + n // world) + 4 * world),
            "peer_bytes_pulled_per_rank_per_step": int(4 * mbar * n / world * (world - 1) / world),
            "setup_s": build_s, "parity": parity,
            "stage_ms_rank0": {k: v / max(1, stage_n) for k, v in stages.items()}}


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    libs = [os.path.join(ROOT, "chanamq_b200", "libchanagpu.so"), os.path.join(ROOT, "chanamq_b200", "libchqgen.so"),
            os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in libs):  # normally prebuilt in-tree by __graft_entry__.build()
        if rank == 0:
            import __graft_entry__ as ge

            ge.build()
        else:
            while not all(os.path.exists(p) for p in libs):
                time.sleep(0.5)
            time.sleep(1.0)
    if a.impl == "reference":
        run_reference(a, rank)
        return
    run_ours(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
