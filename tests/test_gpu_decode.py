"""GPU parity tests of the decode half of the path: chq_decode (k_frame_scan + k_publish_decode) through the C ABI
against the CPU oracle (oracle/amqp_decode.cpp, the restated FrameParser / CommandAssembler / property readers).
Byte work: the records, consumed counts and error codes must be identical."""
import random

import numpy as np
import pytest

from chanamq_b200 import CHQ_MODE_SPEC, CHQ_TOPIC, GpuMatcher
from chanamq_b200 import decode as gd
from chanamq_b200.workload import TopicWorkload, amqp_stream, queue_names_packed
from oracle import pyoracle as po
from tests import amqp_build as ab
from tests.test_oracle_decode import check_vector, vectors

pytestmark = pytest.mark.gpu


def same(data, seg, what=""):
    data = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    seg = np.array(seg, dtype=np.uint64)
    got = gd.decode(data, seg)
    recs, consumed, err, err_off = po.decode(data, seg)
    assert np.array_equal(got.seg_err, err), (what, got.seg_err[:8], err[:8])
    assert np.array_equal(got.seg_consumed, consumed), what
    assert np.array_equal(got.seg_err_off, err_off), what
    assert len(got.recs) == len(recs), (what, len(got.recs), len(recs))
    assert got.recs.tobytes() == recs.tobytes(), what
    return got


def test_reference_vectors_through_the_c_abi():
    for name, (data, want) in vectors().items():
        got = same(data, [0, len(data)], name)
        check_vector(name, data, want, got.recs, got.seg_consumed, got.seg_err, got.seg_err_off)
    # all vectors as segments of ONE batch
    blobs = [d for d, _ in vectors().values()]
    seg = np.cumsum([0] + [len(b) for b in blobs])
    same(b"".join(blobs), seg, "all vectors in one batch")


def test_empty_inputs():
    got = gd.decode(np.zeros(0, np.uint8), np.array([0, 0, 0], np.uint64))
    assert len(got.recs) == 0 and got.seg_consumed.tolist() == [0, 0] and got.seg_err.tolist() == [0, 0]


@pytest.mark.parametrize("body,n_seg", [(0, 7), (100, 64), (4096, 256), (4096, 1)])
def test_generated_streams(body, n_seg):
    w = TopicWorkload(5)
    data, seg, n_msgs = amqp_stream(w, 24_000_000 if n_seg > 1 else 3_000_000, n_seg, body_bytes=body, stream=body + n_seg)
    got = same(data, seg, "body %d" % body)
    assert len(got.recs) == n_msgs and not got.seg_err.any()
    # ragged: cut every segment short at a random place -> the tail command is left for the next call
    rng = random.Random(n_seg)
    cut = seg.copy()
    ends = [int(seg[i]) + rng.randrange(0, int(seg[i + 1] - seg[i]) + 1) for i in range(n_seg)]
    pieces = [data[int(seg[i]):ends[i]] for i in range(n_seg)]
    cut = np.cumsum([0] + [len(p) for p in pieces]).astype(np.uint64)
    same(np.concatenate(pieces) if pieces else data[:0], cut, "ragged")


def test_corrupted_streams_stop_at_the_same_place_with_the_same_code():
    w = TopicWorkload(5)
    data, seg, _ = amqp_stream(w, 4_000_000, 128, body_bytes=200, stream=77)
    rng = random.Random(3)
    bad = data.copy()
    for s in range(0, 128, 2):  # corrupt every other segment once, anywhere
        lo, hi = int(seg[s]), int(seg[s + 1])
        pos = rng.randrange(lo, hi)
        bad[pos] = (int(bad[pos]) + rng.randrange(1, 256)) & 0xFF
    got = same(bad, seg, "fuzz")
    assert (got.seg_err != 0).sum() > 10  # many corruptions land in frame headers / end markers / arguments


def test_decode_then_route_on_device():
    """config 5 in small: raw AMQP bytes -> records -> routing keys -> queue lists, all on the device"""
    torch = pytest.importorskip("torch")
    w = TopicWorkload(2)
    pb, poff, q, d = w.patterns(50_000, 50_000, hash_tail=True)
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC)
    o = po.Oracle(po.TOPIC, po.SPEC, inplace=True)
    m.bind_many(pb, poff, qb, qoff, d)
    o.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    data, seg, n_msgs = amqp_stream(w, 16_000_000, 64, body_bytes=256, stream=9)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d_bytes = torch.from_numpy(data.copy()).to(dev)
    d_seg = torch.from_numpy(seg.view(np.int64).copy()).to(dev)
    cap = n_msgs + 1024
    d_recs = torch.zeros(cap * 8, dtype=torch.int64, device=dev)
    d_misc = torch.zeros(4, dtype=torch.int64, device=dev)  # n_recs, status, n_pub
    d_cons = torch.zeros(64, dtype=torch.int64, device=dev)
    d_err = torch.zeros(64, dtype=torch.int32, device=dev)
    d_eoff = torch.zeros(64, dtype=torch.int64, device=dev)
    gd.decode_device(d_bytes.data_ptr(), len(data), d_seg.data_ptr(), 64, d_recs.data_ptr(), cap, d_misc.data_ptr(),
                     d_cons.data_ptr(), d_err.data_ptr(), d_eoff.data_ptr(), d_misc.data_ptr() + 8, st)
    d_koff = torch.zeros(cap, dtype=torch.int32, device=dev)
    d_klen = torch.zeros(cap, dtype=torch.int32, device=dev)
    d_ridx = torch.zeros(cap, dtype=torch.int32, device=dev)
    gd.publish_keys_device(d_recs.data_ptr(), d_misc.data_ptr(), cap, d_koff.data_ptr(), d_klen.data_ptr(), d_ridx.data_ptr(),
                           d_misc.data_ptr() + 16, st)
    torch.cuda.synchronize()
    n_recs, status, n_pub = int(d_misc[0]), int(d_misc[1]) & 0xFFFFFFFF, int(d_misc[2]) & 0xFFFFFFFF
    assert n_recs == n_msgs and status == 0 and n_pub == n_msgs
    ids_cap = 8_000_000
    d_offsets = torch.zeros(n_pub + 1, dtype=torch.int32, device=dev)
    d_ids = torch.zeros(ids_cap, dtype=torch.int32, device=dev)
    d_dur = torch.zeros(n_pub, dtype=torch.uint8, device=dev)
    d_kerr = torch.zeros(n_pub, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    m.route_keys_device(d_bytes.data_ptr(), d_koff.data_ptr(), d_klen.data_ptr(), n_pub, d_offsets.data_ptr(), d_ids.data_ptr(),
                        ids_cap, d_dur.data_ptr(), d_kerr.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
    torch.cuda.synchronize()
    assert (int(d_tot[1]) & 0xFFFFFFFF) == 0
    # oracle: decode on the CPU, route the extracted keys in the order the device compacted them
    recs, *_ = po.decode(data, seg)
    ridx = d_ridx[:n_pub].cpu().numpy().view(np.uint32)
    assert sorted(ridx.tolist()) == list(range(n_msgs))
    keys = [bytes(data[recs[i]["key_off"]:recs[i]["key_off"] + recs[i]["key_len"]]) for i in ridx]
    kblob, koff = po.pack_keys(keys)
    o_off, o_ids, o_dur, o_err = o.route_packed(kblob, koff, 4)
    assert np.array_equal(d_offsets.cpu().numpy().view(np.uint32), o_off)
    assert np.array_equal(d_ids[: len(o_ids)].cpu().numpy().view(np.uint32), o_ids)
    assert np.array_equal(d_dur.cpu().numpy(), o_dur) and len(o_ids) > 0


@pytest.mark.parametrize("chunk", [64, 300, 4096, 0])
def test_chunked_scan_of_long_segments_is_exact(chunk, monkeypatch):
    """one segment walked by many threads (CHQ_DECODE_CHUNK bytes each; 0 = the one-thread scan): bodies that
    contain whole AMQP messages, heartbeats and frame look-alikes (wrong guesses), frames longer than a chunk
    (chunks without any command start), fragmented bodies, errors and ragged tails"""
    monkeypatch.setenv("CHQ_DECODE_CHUNK", str(chunk))
    rng = random.Random(900 + chunk)
    inner = ab.message("ex", "in.ner.key", b"x" * 40, delivery_mode=2) + ab.heartbeat()

    def body(n):
        kind = rng.randrange(5)
        if kind == 0:
            return bytes(rng.randrange(256) for _ in range(n))
        if kind == 1:   # a message inside the body, repeated: every guess inside it is wrong
            return (inner * (n // len(inner) + 1))[:n]
        if kind == 2:   # bytes that start frames but do not end them
            return (b"\x01\x00\x01\x00\x00\x00\x05" * (n // 7 + 1))[:n]
        if kind == 3:
            return (ab.heartbeat() * (n // 8 + 1))[:n]
        return b"\xce" * n

    def stream(n_msgs):
        out = []
        for i in range(n_msgs):
            n = rng.choice([0, 1, 7, 50, 300, 1000, 5000])
            props = {}
            if rng.random() < 0.3:
                props = dict(delivery_mode=rng.choice([1, 2]), expiration=str(rng.randrange(10**6)),
                             headers=[("h", "S", b"v" * rng.randrange(30)), ("n", "I", i)])
            out.append(ab.message("ex", "k%d.w" % i, body(n), frame_max=rng.choice([None, 64, 1024]), **props))
            if rng.random() < 0.1:
                out.append(ab.heartbeat())
            if rng.random() < 0.05:
                out.append(ab.other_method(60, 80, b"\x00" * 9))
        return b"".join(out)

    segs = [stream(rng.choice([1, 30, 120])) for _ in range(6)]
    segs.append(segs[0][: len(segs[0]) // 2])                    # ragged tail
    segs.append(b"")                                              # empty segment
    broken = bytearray(stream(60))
    broken[len(broken) * 2 // 3] ^= 0x55                          # an error somewhere in the last third
    segs.append(bytes(broken))
    seg = np.cumsum([0] + [len(x) for x in segs])
    got = same(b"".join(segs), seg, "chunk %d" % chunk)
    assert len(got.recs) > 300
    same(segs[1] + segs[2] + segs[3], [0, len(segs[1]) + len(segs[2]) + len(segs[3])], "one long segment")
