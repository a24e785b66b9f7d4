"""Multi-GPU check of the binding-slice path, one rank per GPU over NCCL (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/run_sharded_nccl.py

Every rank keeps slice rank::G of a 200 k-binding table, routes the same 100 k keys, all-gathers, merges, and
compares the merged result with the CPU oracle over the whole table.  Prints one line per rank."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from chanamq_b200 import CHQ_MODE_SPEC, CHQ_TOPIC
    from chanamq_b200.sharded import ShardedGpuMatcher
    from chanamq_b200.workload import TopicWorkload, queue_names_packed
    from oracle import pyoracle as po

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = TopicWorkload(4)
    pb, poff, q, d = w.patterns(200_000, 50_000, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    sm = ShardedGpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC, rank, world, local)
    sm.load(pb, poff, qb, qoff, d)
    blob, off = w.keys(100_000, stream=31)
    n = len(off) - 1
    dev = torch.device("cuda", local)
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    out_off, out_ids, out_dur, err = sm.route_device(d_blob, d_off, n, slice_cap=4_000_000, out_cap=8_000_000)
    o = po.Oracle(po.TOPIC, po.SPEC, inplace=True)
    o.bind_many(pb, poff, qb, qoff, d)
    w_off, w_ids, w_dur, _ = o.route_packed(blob, off, 4)
    ok = (np.array_equal(out_off.cpu().numpy().view(np.uint32), w_off) and
          np.array_equal(out_ids.cpu().numpy().view(np.uint32), w_ids) and np.array_equal(out_dur.cpu().numpy(), w_dur))
    print("rank %d/%d sharded route %s: %d keys -> %d ids" % (rank, world, "OK" if ok else "MISMATCH", n, len(w_ids)), flush=True)
    # the same with the merge sharded by message range (all-to-all of the ids)
    m0, m1, o_off, o_ids, o_dur = sm.route_device_owned(d_blob, d_off, n, slice_cap=4_000_000, out_cap=8_000_000)
    a, b = int(w_off[m0]), int(w_off[m1])
    ok2 = (np.array_equal(o_off.cpu().numpy().view(np.uint32), w_off[m0:m1 + 1] - w_off[m0]) and
           np.array_equal(o_ids.cpu().numpy().view(np.uint32), w_ids[a:b]) and np.array_equal(o_dur.cpu().numpy(), w_dur[m0:m1]))
    print("rank %d/%d owned-range route %s: messages [%d, %d) -> %d ids" % (rank, world, "OK" if ok2 else "MISMATCH", m0, m1, b - a),
          flush=True)
    ok = ok and ok2
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
