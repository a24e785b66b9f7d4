#!/usr/bin/env python
"""Secondary benchmark: BASELINE.json configs[3] as written — the 1 M-binding table SLICED across the GPUs of one
box (binding i on GPU i mod G), every GPU routes ALL keys of a batch against its slice, the per-slice queue lists
are all-gathered over NCCL/NVLink and merged per message on every GPU (chanamq_b200/sharded.py,
chq_merge_slices_device).  bench.py --gpus N measures the other sharding (keys; no collective), which is what the
product recommends whenever the table fits one GPU; this script puts a measured number behind that recommendation.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P bench_sharded.py

Rank 0 prints one JSON line: messages/s of the whole job (one batch is routed ONCE by the G GPUs together),
device time per step as the max over ranks, and the split local route / all-gather / merge."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bindings", type=int, default=1_000_000)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--mode", default="compat", choices=["compat", "spec"])
    ap.add_argument("--merge", default="owned", choices=["owned", "replicated"],
                    help="owned: ids by all-to-all, each rank merges its message range; replicated: all-gather, every rank merges all")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist

    from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC
    from chanamq_b200.sharded import ShardedGpuMatcher
    from chanamq_b200.workload import TopicWorkload, queue_names_packed

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    w = TopicWorkload(4)
    pb, poff, q, d = w.patterns(a.bindings, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    sm = ShardedGpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT if a.mode == "compat" else CHQ_MODE_SPEC, rank, world, local)
    sm.load(pb, poff, qb, qoff, d)
    n = a.batch
    blob, off = w.keys(n, stream=0)            # the SAME batch on every rank
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    slice_cap = int(n * 160 / world * 1.25) + 1024
    out_cap = n * 160
    def step():
        if a.merge == "owned":   # ids by all-to-all, every rank merges the message range it owns
            _, _, _, ids, _ = sm.route_device_owned(d_blob, d_off, n, slice_cap, out_cap // world + 1024)
        else:                    # ids by all-gather, every rank merges everything
            _, ids, _, _ = sm.route_device(d_blob, d_off, n, slice_cap, out_cap)
        return ids

    for _ in range(a.warmup):
        out_ids = step()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / a.steps, float(len(out_ids))], dtype=torch.float64, device=dev)
    tmax = t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    ms = float(tmax[0])
    total = int(t[1]) if a.merge == "owned" else len(out_ids)   # owned: the ranks hold disjoint ranges
    chk = int(out_ids[: min(len(out_ids), 1 << 20)].to(torch.int64).sum())
    stages = None
    if a.merge == "owned":   # one more step with CUDA-event marks between the stages (rank 0's view)
        sm.trace = True
        step()
        stages = {k: round(v, 3) for k, v in sm.last_stages.items()}
    if rank == 0:
        print(json.dumps({
            "metric": "routed msgs/sec (topic, %d bindings sliced over %d GPUs, %s)" % (
                a.bindings, world, "all-to-all of the queue lists, merge sharded by message range" if a.merge == "owned"
                else "NCCL all-gather of queue lists + merge on every GPU"),
            "merge": a.merge, "stages_ms_rank0": stages,
            "value": n / (ms * 1e-3), "unit": "msgs/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong", "dtype": "u32", "data": "synthetic",
            "config": {"workload": "T(seed 4), %d bindings, binding i on GPU i mod %d, one batch of %d keys routed by all GPUs "
                                   "together, %s semantics" % (a.bindings, world, n, a.mode),
                       "mean_queues_per_message": total / n, "slice_cap_ids": slice_cap,
                       "all_gather_bytes_per_rank_per_step": 4 * (n + 1) + 4 * slice_cap + n, "checksum": chk,
                       "timing": "CUDA events around K calls of ShardedGpuMatcher.route_device (local route, three "
                                 "all_gather_into_tensor, chq_merge_slices_device, status read-back), max over ranks"},
        }))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
