"""Binding-slice sharding.  The N > 1 host logic — who owns which binding, global queue
numbering, the all-gather layout, the per-message union — runs here on CPU with world_size 2 over gloo, each
rank matching its slice with the CPU oracle; the GPU tests emulate G slices on one device through the C ABI
(chq_declare_queues + chq_route_device + chq_merge_slices_device) and compare with the unsharded oracle."""
import os
import socket

import numpy as np
import pytest

from chanamq_b200.sharded import merge_slices_host, slice_bindings, slice_of
from chanamq_b200.workload import TopicWorkload, queue_names_packed
from oracle import pyoracle as po


def _table(p, nq, seed=2):
    w = TopicWorkload(seed)
    pb, poff, q, d = w.patterns(p, nq, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    return w, (pb, poff, qb, qoff, d)


def _oracle(binds, mode=po.SPEC, declare=None):
    o = po.Oracle(po.TOPIC, mode, inplace=True)
    assert o.bind_many(*binds) == 0
    return o


def test_slicing_is_a_partition():
    w, binds = _table(1001, 250)
    seen = []
    for r in range(4):
        kb, ko, qb, qo, du = slice_bindings(*binds, r, 4)
        keys = [bytes(kb[ko[i]:ko[i + 1]]) for i in range(len(ko) - 1)]
        want = [bytes(binds[0][binds[1][i]:binds[1][i + 1]]) for i in range(r, 1001, 4)]
        assert keys == want and len(du) == len(want)
        assert [bytes(qb[qo[i]:qo[i + 1]]) for i in range(3)] == [b"q%07d" % ((r + 4 * i) % 250) for i in range(3)]
        seen += list(range(r, 1001, 4))
        assert all(slice_of(i, 4) == r for i in range(r, 1001, 4))
    assert sorted(seen) == list(range(1001))


def _rank_main(rank, world, port, q):
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w, binds = _table(6000, 1500)
        names_all = sorted({bytes(binds[2][binds[3][i]:binds[3][i + 1]]) for i in range(6000)})
        # this rank's slice, matched by the CPU oracle (the product has no CPU matcher); ids are made global
        # the way chq_declare_queues does: rank among ALL queue names of the exchange
        mine = _oracle(slice_bindings(*binds, rank, world))
        local_names = mine.queue_names()
        to_global = np.array([names_all.index(nm) for nm in local_names], dtype=np.uint32)
        blob, off = w.keys(4000, stream=11)  # every rank routes the SAME keys
        n = len(off) - 1
        o_off, o_ids, o_dur, _ = mine.route_packed(blob, off)
        g_ids = to_global[o_ids] if len(o_ids) else o_ids
        # all-gather with padding to a common capacity, as ShardedGpuMatcher.route_device lays it out
        cap = 1 << 16
        t_off = torch.from_numpy(o_off.view(np.int32).copy())
        t_ids = torch.zeros(cap, dtype=torch.int32)
        t_ids[: len(g_ids)] = torch.from_numpy(g_ids.view(np.int32).copy())
        t_dur = torch.from_numpy(o_dur.copy())
        all_off = [torch.empty_like(t_off) for _ in range(world)]
        all_ids = [torch.empty_like(t_ids) for _ in range(world)]
        all_dur = [torch.empty_like(t_dur) for _ in range(world)]
        dist.all_gather(all_off, t_off)
        dist.all_gather(all_ids, t_ids)
        dist.all_gather(all_dur, t_dur)
        offs = np.stack([a.numpy().view(np.uint32) for a in all_off])
        ids = [a.numpy().view(np.uint32) for a in all_ids]
        dur = np.stack([a.numpy() for a in all_dur])
        merged = merge_slices_host(offs, ids, dur)
        # the unsharded oracle over the whole table is the truth
        whole = _oracle(binds)
        assert whole.queue_names() == names_all
        w_off, w_ids, w_dur, _ = whole.route_packed(blob, off)
        ok = (np.array_equal(merged.offsets, w_off) and np.array_equal(merged.queue_ids, w_ids) and
              np.array_equal(merged.any_durable, w_dur))
        q.put((rank, bool(ok), int(len(w_ids)), n))
    finally:
        dist.destroy_process_group()


def test_two_ranks_over_gloo_merge_to_the_unsharded_result():
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(r[0] for r in res) == [0, 1] and all(r[1] for r in res) and res[0][2] > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("g,p,nq", [(2, 30_000, 30_000), (8, 60_000, 15_000), (4, 200_000, 50_000), (4, 200_000, 200_000),
                                    (2, 400_000, 400_000), (8, 400_000, 400_000)])
def test_slices_on_one_gpu_merge_to_the_unsharded_oracle(g, p, nq):
    """G slice tables on one device (what G ranks would hold), routed one after the other, merged by
    chq_merge_slices_device: must equal the oracle over the whole table (4 bindings per queue in two of the
    cases: the same queue then sits in several slices and the union has to de-duplicate)."""
    w, binds = _table(p, nq)
    blob, off = w.keys(50_000, stream=21)
    # one queue per binding (nq == p): no queue sits in two slices, the merge may skip the de-duplication
    _merge_on_one_gpu(g, binds, blob, off, 1 if nq == p else 0)


@pytest.mark.gpu
@pytest.mark.parametrize("g", [2, 3, 8])
def test_slices_with_long_results_merge_to_the_unsharded_oracle(g):
    """Results around and beyond what warp_kway stages (1280 ids): 1200 and 4000 queues behind one pattern each, so a
    message's ids are G runs of 150 ... 2000 — the shared-memory merge at its upper end, and what takes over from it
    (two slices: warp_merge2; more: the long-list union)."""
    pats, queues = [], []
    for i in range(4000):
        pats.append("big.#"); queues.append("b%05d" % ((i * 7) % 4000))
    for i in range(1200):
        pats.append("mid.*"); queues.append("m%05d" % ((i * 11) % 1200))
    for i in range(1290):
        pats.append("edge.*.x"); queues.append("e%05d" % i)
    for i in range(40):
        pats.append("#.few"); queues.append("f%05d" % i)
    order = np.random.RandomState(5).permutation(len(pats))  # which slice a binding lands in: i mod G of this order
    pb, poff = _pack([pats[i] for i in order])
    qb, qoff = _pack([queues[i] for i in order])
    binds = (pb, poff, qb, qoff, (order % 3 == 0).astype(np.uint8))
    keys = ["big.a", "mid.z", "edge.k.x", "big", "mid", "mid.a.b", "big.few", "mid.few", "edge.q.x", "none", "a.few", "big.a.b.c"] * 40
    blob, off = _pack(keys)
    _merge_on_one_gpu(g, binds, blob, off, 1)


def _pack(strings):
    bs = [s.encode() for s in strings]
    off = np.zeros(len(bs) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(b) for b in bs])
    return np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8).copy(), off


def _merge_on_one_gpu(g, binds, blob, off, flags):
    torch = pytest.importorskip("torch")
    from chanamq_b200 import CHQ_MODE_SPEC, CHQ_TOPIC, GpuMatcher
    from chanamq_b200._lib import check, load_library

    L = load_library()
    whole = _oracle(binds)
    n = len(off) - 1
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    cap = 4_000_000
    all_off = torch.empty(g * (n + 1), dtype=torch.int32, device=dev)
    all_ids = torch.zeros(g * cap, dtype=torch.int32, device=dev)
    all_dur = torch.empty(g * n, dtype=torch.uint8, device=dev)
    err = torch.empty(n, dtype=torch.uint8, device=dev)
    tot = torch.zeros(2 * (g + 1), dtype=torch.int64, device=dev)
    qb_all, qo_all = np.ascontiguousarray(binds[2]), np.ascontiguousarray(binds[3])
    tables = []
    for r in range(g):
        m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
        m.bind_many(*slice_bindings(*binds, r, g))
        check(L.chq_declare_queues(m.h, qb_all.ctypes.data, qo_all.ctypes.data, len(qo_all) - 1))
        m.commit()
        assert m.queue_names() == whole.queue_names()
        m.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, all_off.data_ptr() + 4 * r * (n + 1), all_ids.data_ptr() + 4 * r * cap,
                       cap, all_dur.data_ptr() + r * n, err.data_ptr(), tot.data_ptr() + 16 * r, tot.data_ptr() + 16 * r + 8, st)
        tables.append(m)
    out_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
    out_ids = torch.empty(6_000_000, dtype=torch.int32, device=dev)
    out_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    check(L.chq_merge_slices_device2(0, all_off.data_ptr(), all_ids.data_ptr(), cap, all_dur.data_ptr(), g, n, out_off.data_ptr(),
                                     out_ids.data_ptr(), 6_000_000, out_dur.data_ptr(), tot.data_ptr() + 16 * g,
                                     tot.data_ptr() + 16 * g + 8, flags, st))
    torch.cuda.synchronize()
    t = tot.cpu().numpy()
    assert all((int(t[2 * r + 1]) & 0xFFFFFFFF) == 0 for r in range(g + 1))
    w_off, w_ids, w_dur, _ = whole.route_packed(blob, off, 4)
    assert int(t[2 * g]) == len(w_ids)
    assert np.array_equal(out_off.cpu().numpy().view(np.uint32), w_off)
    assert np.array_equal(out_ids[: len(w_ids)].cpu().numpy().view(np.uint32), w_ids)
    assert np.array_equal(out_dur.cpu().numpy(), w_dur)
