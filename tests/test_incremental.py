"""Incremental commit: single binds / unbinds applied in place to a compiled table instead of a
recompile of the whole binding set.  The reference's subscribe / unsubscribe are O(depth) and visible at once
(QueueMatcher.scala:385-401, called per Queue.Bind / Queue.Unbind at ExchangeEntity.scala:333-409).
CPU part: the decision logic on a host-only table (which changes go in place, which fall back).  GPU part: random
bind / unbind / route steps through the C ABI against the oracle, compared as sets of queue NAMES per message (a queue
name first seen by an in-place bind takes the next free id, so ids are name ranks again only after a full commit)."""
import random
import time

import numpy as np
import pytest

from chanamq_b200 import CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC, CHQ_DIRECT, GpuMatcher
from oracle import pyoracle as po
from tests.helpers import pack


def test_which_changes_go_in_place_host_only():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC, device=-1)
    for i in range(200):
        m.bind("a.b%d.*" % i, "q%d" % (i % 7))
        m.bind("x.%d" % i, "q%d" % (i % 5))
    m.commit()
    st = m.stats()
    assert st["full_commits"] == 1 and st["incremental_commits"] == 0
    g0 = st["groups"]
    m.bind("a.b1.*", "q_new")          # known pattern, new queue name
    m.bind("a.zz.*", "q1")             # new pattern and new word, known shape
    m.bind("x.7", "q2")                # one more queue on a known pattern
    m.unbind("x.3", "q3")
    m.commit()
    st = m.stats()
    assert st["full_commits"] == 1 and st["incremental_commits"] == 1 and st["groups"] == g0 + 1
    assert st["queues"] == 8 and m.queue_names()[-1] == b"q_new"     # appended, not ranked
    m.bind("a.b.c.d.e", "q1")          # a shape the table has not seen: full compile
    m.commit()
    st = m.stats()
    assert st["full_commits"] == 2 and m.queue_names() == sorted(m.queue_names())   # ranked by name again
    m.bind("a.#.b", "q1")              # '#' inside a pattern widens the word rows: full compile
    m.commit()
    assert m.stats()["full_commits"] == 3 and m.stats()["wstride"] == 256
    m.bind("a.b1.", "q1")              # not in canonical form (trailing dot): full compile
    m.commit()
    assert m.stats()["full_commits"] == 4
    m.bind("a.b2.*", "q_other")
    m.commit_full()                    # on request
    assert m.stats()["full_commits"] == 5


def _names_per_message(res, names):
    off = res.offsets
    return [sorted(names[j] for j in res.queue_ids[off[i]:off[i + 1]]) for i in range(len(off) - 1)]


def _oracle_names(o, blob, off):
    o_off, o_ids, o_dur, _ = o.route_packed(blob, off, 1)
    names = o.queue_names()
    return [sorted(names[j] for j in o_ids[o_off[i]:o_off[i + 1]]) for i in range(len(o_off) - 1)], o_dur


@pytest.mark.gpu
@pytest.mark.parametrize("mode,xtype", [(CHQ_MODE_COMPAT, CHQ_TOPIC), (CHQ_MODE_SPEC, CHQ_TOPIC), (CHQ_MODE_COMPAT, CHQ_DIRECT)])
def test_random_bind_unbind_route_steps_match_the_oracle(mode, xtype):
    rng = random.Random(40 + mode + 10 * xtype)
    alpha = ["a", "b", "c", "dd", "eeeeeeeeee", "f"]
    wild = ["*", "*", "#"]
    omode = po.SPEC if mode == CHQ_MODE_SPEC else po.COMPAT
    m = GpuMatcher(xtype, mode)
    o = po.Oracle(po.TOPIC if xtype == CHQ_TOPIC else po.DIRECT, omode)

    def topic(pw):
        n = rng.randint(1, 4)
        return ".".join(rng.choice(wild) if rng.random() < pw else rng.choice(alpha) for _ in range(n))

    live = set()
    for _ in range(600):   # the table the broker starts from (restart: replay of the stored binds)
        k, q, d = topic(0.3), "q%03d" % rng.randrange(40), rng.random() < 0.3
        if m.try_bind(k, q, d) == 0:
            assert o.bind(k, q, d) == 0
            live.add((k, q, d))
    m.commit()
    keys = [topic(0.0) for _ in range(400)] + ["a", "a.b", "*", "#", "a.#"]
    blob, off = pack(keys)
    for step in range(60):
        for _ in range(rng.randint(1, 4)):
            if live and rng.random() < 0.4:
                k, q, d = rng.choice(sorted(live))
                # (sequences that avoid the reference's tombstone defect, SURVEY A.6, are not needed: the oracle is
                # rebuilt from the live set below when an unbind happened)
                m.unbind(k, q, d)
                live.discard((k, q, d))
                o = None
            else:
                k, q, d = topic(0.3), "q%03d" % rng.randrange(60), rng.random() < 0.3   # new names appear too
                if m.try_bind(k, q, d) == 0:
                    live.add((k, q, d))
                    if o is not None:
                        assert o.bind(k, q, d) == 0
        if o is None:
            o = po.Oracle(po.TOPIC if xtype == CHQ_TOPIC else po.DIRECT, omode)
            for k, q, d in sorted(live):
                assert o.bind(k, q, d) == 0
        m.commit()
        got = m.route_packed(blob, off)
        want, want_dur = _oracle_names(o, blob, off)
        assert _names_per_message(got, m.queue_names()) == want, "step %d" % step
        assert np.array_equal(got.any_durable, want_dur), "step %d" % step
        # ascending ids inside every message, whatever the numbering
        for i in range(len(keys)):
            ids = got.queue_ids[got.offsets[i]:got.offsets[i + 1]]
            assert np.all(np.diff(ids.astype(np.int64)) > 0)
    st = m.stats()
    assert st["incremental_commits"] > 20, st
    # and a full commit brings the same sets back with name-ranked ids
    m.commit_full()
    got = m.route_packed(blob, off)
    assert _names_per_message(got, m.queue_names()) == want and m.queue_names() == sorted(m.queue_names())


@pytest.mark.gpu
def test_one_bind_on_the_1m_table_is_cheap():
    """the point of it: a Queue.Bind on the benchmark's table must not stall routing for seconds"""
    from chanamq_b200.workload import TopicWorkload, queue_names_packed

    w = TopicWorkload(4)
    pb, poff, q, d = w.patterns(1_000_000, 1_000_000, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    m.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    key = bytes(pb[poff[5]:poff[6]]).decode()
    blob, off = pack([key.replace("*", "zz")] * 4)
    before = _names_per_message(m.route_packed(blob, off), m.queue_names())
    t0 = time.perf_counter()
    m.bind(key, "a-brand-new-queue")
    m.commit()
    dt = time.perf_counter() - t0
    after = _names_per_message(m.route_packed(blob, off), m.queue_names())
    assert m.stats()["incremental_commits"] == 1
    assert dt < 0.05, "bind + commit took %.3f s" % dt
    if before[0]:   # the key matched its own pattern: the new queue is there now
        assert after[0] == sorted(before[0] + [b"a-brand-new-queue"])
