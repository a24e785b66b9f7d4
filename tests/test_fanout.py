"""The step after the match: per-queue message lists in publish order and the per-message
flags of ExchangeEntity.publish (chana-mq-server/.../entity/ExchangeEntity.scala:300-307, 319-330).
CPU: the numpy oracle against a literal replay of the reference loop.  GPU: chq_fanout_device /
chq_publish_flags_device through the C ABI against that oracle — index work, so bit-exact."""
import random

import numpy as np
import pytest

from oracle import pyoracle as po


def _replay(offsets, ids, n_queues):
    """ExchangeEntity.scala:305-307, message by message"""
    q = [[] for _ in range(n_queues)]
    for m in range(len(offsets) - 1):
        for e in range(offsets[m], offsets[m + 1]):
            q[ids[e]].append(m)
    off = np.zeros(n_queues + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in q])
    return off, np.array([m for x in q for m in x], dtype=np.uint32)


def _random_csr(rng, n, n_queues, density):
    sets = [sorted(rng.sample(range(n_queues), min(n_queues, rng.choice(density)))) for _ in range(n)]
    off = np.zeros(n + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(s) for s in sets])
    return off, np.array([q for s in sets for q in s], dtype=np.uint32)


def test_oracle_transpose_equals_the_reference_loop():
    rng = random.Random(1)
    for n, nq, dens in [(0, 3, [0]), (1, 1, [1]), (200, 17, [0, 1, 2, 9]), (500, 300, [0, 0, 3, 40])]:
        off, ids = _random_csr(rng, n, nq, dens)
        a, b = po.fanout_transpose(off, ids, nq)
        c, d = _replay(off, ids, nq)
        assert np.array_equal(a, c) and np.array_equal(b, d)


def test_oracle_flags():
    off = np.array([0, 2, 2, 3, 3], dtype=np.uint32)
    ids = np.array([0, 1, 1], dtype=np.uint32)
    dur = np.array([1, 0, 0, 0], dtype=np.uint8)
    pf = np.array([4 | 2, 1 | 2, 2 | 4, 0], dtype=np.uint8)     # persist+immediate, mandatory+immediate, immediate+persist, none
    cnt, fl = po.publish_flags(off, ids, dur, pf, True, np.array([0, 1], dtype=np.uint8))
    assert cnt.tolist() == [2, 0, 1, 0]
    assert fl.tolist() == [1, 2 | 4, 0, 0]      # msg 0 persistent, has a consumer on q1; msg 1 unrouted; msg 2 q1 has a consumer
    _, fl = po.publish_flags(off, ids, dur, pf, False, None)
    assert fl.tolist() == [4, 2 | 4, 4, 0]


def _gpu_transpose(off, ids, n_queues, q_cap=None):
    import torch

    from chanamq_b200 import fanout as gf

    dev = torch.device("cuda:0")
    n = len(off) - 1
    total = int(off[-1])
    cap = total if q_cap is None else q_cap
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    d_ids = torch.from_numpy(np.ascontiguousarray(ids).view(np.int32).copy() if total else np.zeros(1, np.int32)).to(dev)
    d_qoff = torch.full((n_queues + 1,), -1, dtype=torch.int32, device=dev)
    d_qmsg = torch.full((max(cap, 1) + 1,), -7, dtype=torch.int32, device=dev)
    d_st = torch.full((1,), 99, dtype=torch.int32, device=dev)
    gf.fanout_device(d_off.data_ptr(), d_ids.data_ptr(), n, n_queues, d_qoff.data_ptr(), d_qmsg.data_ptr(), cap, d_st.data_ptr(),
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return int(d_st[0]), d_qoff.cpu().numpy().view(np.uint32), d_qmsg.cpu().numpy().view(np.uint32)


@pytest.mark.gpu
@pytest.mark.parametrize("n,nq,dens", [(1, 1, [1]), (33, 5, [0, 1, 5]), (5000, 4096, [0, 1, 3, 60]), (4097, 4097, [2]),
                                       (20000, 9, [0, 4, 9]), (3000, 40000, [0, 0, 500])])
def test_transpose_against_the_oracle(n, nq, dens):
    """ranges of every size class: empty, one entry, registers (<= 32), shared memory (<= 1024), bitmap (longer)"""
    rng = random.Random(n * 31 + nq)
    off, ids = _random_csr(rng, n, nq, dens)
    st, q_off, q_msg = _gpu_transpose(off, ids, nq)
    w_off, w_msg = po.fanout_transpose(off, ids, nq)
    assert st == 0
    assert np.array_equal(q_off, w_off)
    assert np.array_equal(q_msg[: len(w_msg)], w_msg)
    assert q_msg[len(w_msg)] == np.uint32(-7 & 0xFFFFFFFF)   # nothing written past the end


@pytest.mark.gpu
def test_transpose_edge_cases():
    st, q_off, _ = _gpu_transpose(np.zeros(1, np.uint32), np.zeros(0, np.uint32), 5)      # empty batch
    assert st == 0 and q_off.tolist() == [0] * 6
    off = np.array([0, 0, 0, 0], np.uint32)                                                # nothing routed
    st, q_off, _ = _gpu_transpose(off, np.zeros(0, np.uint32), 3)
    assert st == 0 and q_off.tolist() == [0, 0, 0, 0]
    off, ids = np.array([0, 2, 3], np.uint32), np.array([0, 1, 1], np.uint32)
    st, q_off, q_msg = _gpu_transpose(off, ids, 2, q_cap=2)                                # too small a buffer
    assert st == -3 and q_msg[2] == np.uint32(-7 & 0xFFFFFFFF)
    st, _, _ = _gpu_transpose(off, np.array([0, 1, 7], np.uint32), 2)                      # an id outside the table
    assert st == -5


@pytest.mark.gpu
def test_overflow_leaves_the_memory_behind_the_buffer_alone():
    """CHQ_E_OVERFLOW: the ranges of the full counts reach past q_cap; neither the scatter nor the sorts may touch
    what lies there (small DESCENDING sentinels: any sort over them would reorder them)"""
    import torch

    from chanamq_b200 import fanout as gf

    rng = random.Random(77)
    n, nq = 6000, 40
    off, ids = _random_csr(rng, n, nq, [0, 3, 9])
    total = int(off[-1])
    cap, tail = total // 2, 4096
    dev = torch.device("cuda:0")
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    d_ids = torch.from_numpy(np.ascontiguousarray(ids).view(np.int32).copy()).to(dev)
    d_qoff = torch.full((nq + 1,), -1, dtype=torch.int32, device=dev)
    sentinel = np.arange(tail, 0, -1, dtype=np.int32)
    d_qmsg = torch.cat([torch.zeros(cap, dtype=torch.int32), torch.from_numpy(sentinel)]).to(dev)
    d_st = torch.full((1,), 99, dtype=torch.int32, device=dev)
    gf.fanout_device(d_off.data_ptr(), d_ids.data_ptr(), n, nq, d_qoff.data_ptr(), d_qmsg.data_ptr(), cap, d_st.data_ptr(),
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert int(d_st[0]) == -3
    assert np.array_equal(d_qmsg.cpu().numpy()[cap:], sentinel)


@pytest.mark.gpu
def test_routed_batch_to_queue_lists_and_flags():
    """route 2^17 keys against a 100 k-binding table on the device, transpose, flags — all without leaving HBM"""
    import torch

    from chanamq_b200 import CHQ_MODE_SPEC, CHQ_TOPIC, GpuMatcher
    from chanamq_b200 import fanout as gf
    from chanamq_b200.workload import TopicWorkload, queue_names_packed

    w = TopicWorkload(2)
    pb, poff, q, d = w.patterns(100_000, 25_000, hash_tail=True)   # shared queues: dedup upstream
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
    m.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    blob, off = w.keys(1 << 17, stream=4)
    n, nq = len(off) - 1, m.n_queues()
    r = m.route_packed(blob, off)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d_off = torch.from_numpy(r.offsets.view(np.int32).copy()).to(dev)
    d_ids = torch.from_numpy(r.queue_ids.view(np.int32).copy()).to(dev)
    d_dur = torch.from_numpy(r.any_durable.copy()).to(dev)
    total = len(r.queue_ids)
    d_qoff = torch.empty(nq + 1, dtype=torch.int32, device=dev)
    d_qmsg = torch.empty(total, dtype=torch.int32, device=dev)
    d_st = torch.zeros(1, dtype=torch.int32, device=dev)
    rng = np.random.default_rng(5)
    pf = rng.integers(0, 8, n).astype(np.uint8)
    hc = (rng.random(nq) < 0.3).astype(np.uint8)
    d_pf, d_hc = torch.from_numpy(pf).to(dev), torch.from_numpy(hc).to(dev)
    d_cnt = torch.empty(n, dtype=torch.int32, device=dev)
    d_fl = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(2):   # twice: the result does not depend on the order the atomics arrive in
        gf.fanout_device(d_off.data_ptr(), d_ids.data_ptr(), n, nq, d_qoff.data_ptr(), d_qmsg.data_ptr(), total, d_st.data_ptr(), st)
        gf.publish_flags_device(d_off.data_ptr(), d_ids.data_ptr(), d_dur.data_ptr(), d_pf.data_ptr(), n, True, d_hc.data_ptr(), nq,
                                d_cnt.data_ptr(), d_fl.data_ptr(), st)
        torch.cuda.synchronize()
        w_off, w_msg = po.fanout_transpose(r.offsets, r.queue_ids, nq)
        assert int(d_st[0]) == 0
        assert np.array_equal(d_qoff.cpu().numpy().view(np.uint32), w_off)
        assert np.array_equal(d_qmsg.cpu().numpy().view(np.uint32), w_msg)
        w_cnt, w_fl = po.publish_flags(r.offsets, r.queue_ids, r.any_durable, pf, True, hc)
        assert np.array_equal(d_cnt.cpu().numpy().view(np.uint32), w_cnt) and np.array_equal(d_fl.cpu().numpy(), w_fl)
    # size-independent properties: a permutation of the message indices, ascending inside every queue
    assert int(w_off[-1]) == total and w_msg.max() < n
