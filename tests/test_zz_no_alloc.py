"""chq_route_device_static — the entry point for callers that may not allocate device memory (a
routing thread inside a latency budget, a process whose memory all belongs to a pool, a deployment
that forbids allocation after start-up).  It runs the same kernels as chq_route_device with its
scratch in a fixed array inside the library, so the check is simple: through the C ABI, on the same
device buffers, it must give exactly what chq_route_device gives AND what the CPU oracle gives (integer
work: equality) — on two calls in a row, because the array keeps the previous call's contents — and it
must refuse, loudly and without launching anything, a batch larger than the array was sized for and a
table whose routing needs scratch the array does not carry (headers x-match).

Collected last on purpose (file name): a failure here cannot hide another test's result under -x."""
import numpy as np
import pytest

from chanamq_b200 import CHQ_HEADERS, CHQ_MODE_SPEC, CHQ_TOPIC, ChqError, GpuMatcher
from chanamq_b200 import _lib
from chanamq_b200.workload import TopicWorkload, queue_names_packed

pytestmark = pytest.mark.gpu

STATIC_MAX_KEYS = 65536


def _run(torch, m, entry, blob, off, cap):
    dev = torch.device("cuda:0")
    n = len(off) - 1
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    d_offsets = torch.full((n + 1,), -1, dtype=torch.int32, device=dev)
    d_ids = torch.full((cap,), -1, dtype=torch.int32, device=dev)
    d_dur = torch.full((n,), 7, dtype=torch.uint8, device=dev)
    d_err = torch.full((n,), 7, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    entry(d_blob.data_ptr(), d_off.data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), cap, d_dur.data_ptr(),
          d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    total, status = int(d_tot[0].item()), int(d_tot[1].item()) & 0xFFFFFFFF
    return (status, total, d_offsets.cpu().numpy().view(np.uint32), d_ids[:max(total, 0)].cpu().numpy().view(np.uint32),
            d_dur.cpu().numpy(), d_err.cpu().numpy())


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    torch = pytest.importorskip("torch")
    w = TopicWorkload(2)
    pb, poff, q, d = w.patterns(60_000, 15_000, hash_tail=True)  # shared queues: the de-duplicating merge runs too
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
    m.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    from oracle import pyoracle as po
    o = po.Oracle(po.TOPIC, po.SPEC, inplace=True)
    assert o.bind_many(pb, poff, qb, qoff, d) == 0
    cap = 4_000_000
    # second call: fewer keys of another stream, plus keys the first call never saw (errors, long, empty)
    extra = [b".", b"..", b"", b"x" * 300, b"a.b.c.d.e.f.g.h.i.j"]
    for call, (n, stream) in enumerate([(STATIC_MAX_KEYS, 40), (30_001, 41)]):
        blob, off = w.keys(n - (len(extra) if call else 0), stream=stream)
        if call:
            tail = np.frombuffer(b"".join(extra), dtype=np.uint8)
            off = np.concatenate([off, off[-1] + np.cumsum([len(e) for e in extra]).astype(np.uint32)]).astype(np.uint32)
            blob = np.concatenate([blob[: off[-1 - len(extra)]], tail])
        want = _run(torch, m, m.route_device, blob, off, cap)
        got = _run(torch, m, m.route_device_static, blob, off, cap)
        assert want[0] == 0 and got[0] == 0, (want[0], got[0])
        assert got[1] == want[1] and want[1] > 0
        for a, b in zip(want[2:], got[2:]):
            assert np.array_equal(a, b), "call %d" % call
        # and against the CPU oracle: offsets, ids, durable flags, error flags
        o_off, o_ids, o_dur, o_err = o.route_packed(blob, off, 4)
        assert np.array_equal(got[2], o_off) and np.array_equal(got[3], o_ids), "call %d vs oracle" % call
        # (a key longer than a shortstr is flagged by the product, err 2; the oracle takes any string)
        short = np.diff(off.astype(np.int64)) <= 255
        assert np.array_equal(got[4], o_dur) and np.array_equal((got[5] != 0)[short], (o_err != 0)[short]), "call %d vs oracle" % call
    # the slow merge path through the fixed arena (limits forced down), still identical
    m.debug_limits(1, 1)
    blob, off = w.keys(20_000, stream=42)
    want = _run(torch, m, m.route_device, blob, off, cap)
    got = _run(torch, m, m.route_device_static, blob, off, cap)
    assert want[0] == 0 and got[0] == 0
    assert all(np.array_equal(a, b) for a, b in zip(want[2:], got[2:]))
    o_off, o_ids, o_dur, _ = o.route_packed(blob, off, 4)
    assert np.array_equal(got[2], o_off) and np.array_equal(got[3], o_ids) and np.array_equal(got[4], o_dur)
    m.debug_limits(0, 0)


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    torch = pytest.importorskip("torch")
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
    m.bind("a.*", "q")
    dummy = torch.zeros(64, dtype=torch.int64, device="cuda:0").data_ptr()
    with pytest.raises(ChqError) as e:
        m.route_device_static(dummy, dummy, STATIC_MAX_KEYS + 1, dummy, dummy, 1, dummy, dummy, dummy, dummy)
    assert e.value.code == _lib.CHQ_E_ARG and "CHQ_STATIC_MAX_KEYS" in str(e.value)
    m.bind("a.#.b", "q2")  # '#' inside a pattern: 256 word ids are kept per key, so far fewer keys fit
    with pytest.raises(ChqError) as e:
        m.route_device_static(dummy, dummy, 5000, dummy, dummy, 1, dummy, dummy, dummy, dummy)
    assert e.value.code == _lib.CHQ_E_ARG and "CHQ_STATIC_MAX_WORDS" in str(e.value)
    # and a batch that does fit still routes correctly on that table
    keys = [b"a.x", b"a.b", b"a.q.r.b", b"b"]
    from oracle import pyoracle as po
    blob, off = po.pack_keys(keys)
    st, total, offs, ids, dur, err = _run(torch, m, m.route_device_static, blob, off, 64)
    assert st == 0 and offs.tolist() == [0, 1, 3, 4, 4] and ids.tolist() == [0, 0, 1, 1]


def test_static_scratch_entry_point_refuses_a_headers_table():
    """x-match routing needs a per-batch descriptor arena the fixed array does not carry: the call must say so
    instead of treating the header tables as routing keys and matching nothing"""
    torch = pytest.importorskip("torch")
    m = GpuMatcher(CHQ_HEADERS, CHQ_MODE_SPEC)
    # arguments table: "x-match" = S "any", "k" = S "v"
    args = b"\x07x-matchS\x00\x00\x00\x03any" + b"\x01kS\x00\x00\x00\x01v"
    m.bind_args("", "q", False, args)
    m.commit()
    dummy = torch.zeros(64, dtype=torch.int64, device="cuda:0").data_ptr()
    with pytest.raises(ChqError) as e:
        m.route_device_static(dummy, dummy, 1, dummy, dummy, 1, dummy, dummy, dummy, dummy)
    assert e.value.code == _lib.CHQ_E_ARG and "headers" in str(e.value)
