"""Shared by the parity tests: load the same bindings into the CUDA table (through the C ABI) and
into the CPU oracle, route the same keys through both, compare bit for bit."""
import numpy as np

from oracle import pyoracle as po


def pack(strings):
    return po.pack_keys(strings)


def load_both(matcher, oracle, binds):
    """binds: iterable of (key, queue, durable).  Both sides must agree on every bind status."""
    for key, queue, durable in binds:
        rc_o = oracle.bind(key, queue, durable)
        rc_g = matcher.try_bind(key, queue, durable)
        # -1 zero words on both sides; the product additionally rejects >255-byte keys (-2)
        assert (rc_o == 0) == (rc_g == 0) or rc_g == -2, (key, rc_o, rc_g)
        if rc_g == -2 and rc_o == 0:
            oracle.unbind(key, queue, durable)


def assert_same_routing(matcher, oracle, blob, off, threads=1, what=""):
    got = matcher.route_packed(blob, off)
    o_off, o_ids, o_dur, o_err = oracle.route_packed(blob, off, threads)
    assert matcher.queue_names() == oracle.queue_names(), what
    n = len(off) - 1
    assert got.offsets.shape == (n + 1,)
    if not np.array_equal(got.err != 0, o_err != 0):
        bad = np.nonzero((got.err != 0) != (o_err != 0))[0][:5]
        raise AssertionError("%s: err flags differ at %s" % (what, bad))
    if not np.array_equal(got.offsets, o_off):
        i = int(np.nonzero(got.offsets != o_off)[0][0]) - 1
        i = max(i, 0)
        key = bytes(blob[off[i]:off[i + 1]])
        raise AssertionError("%s: message %d key %r: gpu %s, oracle %s" % (
            what, i, key, got.queue_ids[got.offsets[i]:got.offsets[i + 1]][:20], o_ids[o_off[i]:o_off[i + 1]][:20]))
    if not np.array_equal(got.queue_ids, o_ids):
        j = int(np.nonzero(got.queue_ids != o_ids)[0][0])
        i = int(np.searchsorted(o_off, j, side="right")) - 1
        key = bytes(blob[off[i]:off[i + 1]])
        raise AssertionError("%s: message %d key %r: gpu %s, oracle %s" % (
            what, i, key, got.queue_ids[o_off[i]:o_off[i + 1]][:20], o_ids[o_off[i]:o_off[i + 1]][:20]))
    assert np.array_equal(got.any_durable, o_dur), what
    return got


def check_csr_invariants(res, n_queues):
    """size-independent properties of a routing result"""
    off = res.offsets.astype(np.int64)
    assert off[0] == 0 and np.all(np.diff(off) >= 0)
    assert len(res.queue_ids) == off[-1]
    ids = res.queue_ids.astype(np.int64)
    if len(ids):
        assert ids.max() < n_queues
        # strictly ascending inside every message: a set, in canonical order
        d = np.diff(ids)
        starts = off[1:-1][(off[1:-1] > 0) & (off[1:-1] < len(ids))]
        interior = np.ones(len(d), dtype=bool)
        interior[starts - 1] = False
        assert np.all(d[interior] > 0)
    assert np.all(res.offsets[1:][res.err != 0] == res.offsets[:-1][res.err != 0])
