"""What the box can carry from the GPUs to host memory at once: every rank copies `--mb` MB (default: the id array
of one headline batch) from its GPU into pinned host memory, over and over, all ranks together; prints GB/s per rank
and in total — the ceiling of the end-to-end figure at N GPUs.  With --numa each rank first moves itself to its GPU's
NUMA node (chq_numa_bind_to_device), which is what bench.py does.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29544 \
      tools/d2h_probe.py [--numa] [--mb 415] [--iters 20]"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--numa", action="store_true")
ap.add_argument("--mb", type=int, default=415)
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
node = -1
if a.numa:
    from chanamq_b200._lib import load_library

    nd, nc = C.c_int32(-1), C.c_int32(0)
    load_library().chq_numa_bind_to_device(local, C.byref(nd), C.byref(nc))
    node = nd.value
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = a.mb * 1_000_000
src = torch.empty(n, dtype=torch.uint8, device="cuda")
dst = torch.empty(n, dtype=torch.uint8).pin_memory()
for _ in range(2):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.iters):
    dst.copy_(src, non_blocking=True)
e1.record()
torch.cuda.synchronize()
gbs = n * a.iters / (e0.elapsed_time(e1) * 1e-3) / 1e9
t = torch.tensor([gbs], dtype=torch.float64, device="cuda")
allg = [torch.zeros_like(t) for _ in range(world)]
if world > 1:
    dist.all_gather(allg, t)
else:
    allg = [t]
if rank == 0:
    per = [float(x[0]) for x in allg]
    print(json.dumps({"what": "concurrent pinned device-to-host copies", "n_gpus": world, "mb_per_copy": a.mb, "numa_bound": a.numa,
                      "gb_s_per_rank": per, "gb_s_total": sum(per), "affinity_rank0": sorted(os.sched_getaffinity(0))[:4] + ["..."],
                      "numa_node_rank0": node}))
if world > 1:
    dist.destroy_process_group()
