"""Binding-slice sharding as the product does it: chq_shard_* of the C ABI, one process per GPU, NCCL and the peer
reads inside libchanagpu.so (BASELINE.json configs[3]).  Needs two GPUs (skipped otherwise; run it with
`gpurun --gpus 2 -- python -m pytest tests/test_shard_nccl.py -m gpu`).  Every rank routes the same keys against its
slice and must come back with exactly the oracle's result (whole table, one CPU trie) for its own range of messages:
with shared queues (the unions de-duplicate across slices) and with one queue per binding (they need not), over
several batches in a row (the slice buffers alternate), and with capacities too small (every rank must report the
overflow instead of waiting for the others in a collective)."""
import multiprocessing as mp
import os
import sys
import traceback

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rank_main(rank, world, ids, out):
    try:
        sys.path.insert(0, ROOT)
        import torch

        from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC
        from chanamq_b200._lib import CHQ_E_OVERFLOW
        from chanamq_b200.sharded import ShardedExchange
        from chanamq_b200.workload import TopicWorkload, queue_names_packed
        from oracle import pyoracle as po

        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        st = torch.cuda.current_stream().cuda_stream
        report = []
        for case, (mode, omode, p, nq) in enumerate([(CHQ_MODE_SPEC, po.SPEC, 200_000, 49_999), (CHQ_MODE_COMPAT, po.COMPAT, 300_000, 300_000)]):
            w = TopicWorkload(4)
            pb, poff, q, d = w.patterns(p, nq, hash_tail=True)
            qb, qoff = queue_names_packed(q)
            ex = ShardedExchange(ids[case], rank, world, rank, CHQ_TOPIC, mode)
            ex.bind_many(pb, poff, qb, qoff, d)
            ex.commit()
            assert ex.disjoint == (nq == p)
            o = po.Oracle(po.TOPIC, omode, inplace=True)
            assert o.bind_many(pb, poff, qb, qoff, d) == 0
            n = 100_000
            cap = 12_000_000
            d_offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
            d_ids = torch.empty(cap, dtype=torch.int32, device=dev)
            d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
            d_err = torch.empty(n, dtype=torch.uint8, device=dev)
            d_tot = torch.zeros(2, dtype=torch.int64, device=dev)

            def route(stream_no, n_keys):
                blob, off = w.keys(n_keys, stream=stream_no)
                d_blob = torch.from_numpy(blob.copy()).to(dev)
                d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
                m0, m1 = ex.route_device(d_blob.data_ptr(), d_off.data_ptr(), n_keys, d_offsets.data_ptr(), d_ids.data_ptr(), cap,
                                         d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
                torch.cuda.synchronize()
                status = int(d_tot[1].item()) & 0xFFFFFFFF
                status = status - (1 << 32) if status >= (1 << 31) else status
                return blob, off, m0, m1, status

            # capacities too small for one rank's range: every rank says so, nobody hangs
            ex.reserve(n, 12_000_000, 1000)
            _, _, _, _, status = route(30, n)
            assert status == CHQ_E_OVERFLOW, status
            ex.reserve(n, 12_000_000, 8_000_000)
            for b, n_keys in enumerate([n, n - 12_345, n]):
                blob, off, m0, m1, status = route(31 + b, n_keys)
                assert status == 0, status
                assert (m0, m1) == (rank * n_keys // world, (rank + 1) * n_keys // world)
                w_off, w_ids, w_dur, w_err = o.route_packed(blob, off, 4)
                nn = m1 - m0
                g_off = d_offsets[: nn + 1].cpu().numpy().view(np.uint32)
                a, e = int(w_off[m0]), int(w_off[m1])
                assert np.array_equal(g_off, w_off[m0:m1 + 1] - w_off[m0]), "case %d batch %d offsets" % (case, b)
                assert int(d_tot[0].item()) == e - a
                assert np.array_equal(d_ids[: e - a].cpu().numpy().view(np.uint32), w_ids[a:e]), "case %d batch %d ids" % (case, b)
                assert np.array_equal(d_dur[:nn].cpu().numpy(), w_dur[m0:m1])
                assert np.array_equal(d_err[:nn].cpu().numpy() != 0, w_err[m0:m1] != 0)
                report.append((case, b, m0, m1, e - a))
            ex.close()
        out.put((rank, "ok", report))
    except Exception:
        out.put((rank, "fail", traceback.format_exc()))


def test_two_ranks_through_the_c_abi_match_the_oracle():
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from chanamq_b200.sharded import ShardedExchange

    world = 2
    ids = [ShardedExchange.unique_id(), ShardedExchange.unique_id()]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, world, ids, out)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    try:
        for _ in range(world):
            results.append(out.get(timeout=600))
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    for rank, verdict, detail in sorted(results):
        assert verdict == "ok", "rank %d:\n%s" % (rank, detail)
