"""Delivery-side encode: chq_deliver_encode_device against the restated AMQCommand.render
(oracle/pyoracle.py: render_deliver; chana-mq-base/.../model/AMQCommand.scala:30-65).  Byte work: the rendered
streams must be identical.  The publishes are built with the restated property writers of tests/amqp_build.py,
decoded on the device, and every record is delivered a few times with different consumers and frame sizes."""
import random
import struct

import numpy as np
import pytest

from oracle import pyoracle as po
from tests import amqp_build as ab


def _publishes(rng, n):
    """-> (stream bytes, [(exchange, key, header payload, body, channel)])"""
    out, info = [], []
    for i in range(n):
        body = bytes(rng.randrange(256) for _ in range(rng.choice([0, 1, 7, 15, 16, 17, 100, 1000, 5000])))
        props = {}
        if rng.random() < 0.5:
            props["delivery_mode"] = rng.choice([1, 2])
        if rng.random() < 0.3:
            props["expiration"] = str(rng.randrange(10**6))
        if rng.random() < 0.4:
            props["headers"] = [("h%d" % j, "S", b"v" * rng.randrange(20)) for j in range(rng.randrange(0, 5))]
        ch = rng.randrange(1, 2000)
        ex, key = "ex%d" % rng.randrange(3), ".".join("w%d" % rng.randrange(50) for _ in range(rng.randrange(1, 5)))
        hdr_frame = ab.content_header(len(body), channel=ch, **props)
        out.append(ab.message(ex, key, body, channel=ch, frame_max=rng.choice([None, 64, 1000]), **props))
        if rng.random() < 0.1:
            out.append(ab.heartbeat())
        info.append((ex.encode(), key.encode(), hdr_frame[7:-1], body, ch))
    return b"".join(out), info


def test_oracle_render_is_what_the_decoder_reads_back():
    """render_deliver output parses as frames with the right sizes (no GPU: frame arithmetic only)"""
    rng = random.Random(1)
    _, info = _publishes(rng, 30)
    for ex, key, hp, body, ch in info:
        for fm in (0, 9, 64, 4096):
            s = po.render_deliver(ch, b"ctag", 5, False, ex, key, hp, body, fm)
            p, types, got = 0, [], b""
            while p < len(s):
                t, c, n = struct.unpack(">BHI", s[p:p + 7])
                assert c == ch and s[p + 7 + n] == 0xCE
                if t == 3:
                    assert fm == 0 or n <= fm - 8
                    got += s[p + 7:p + 7 + n]
                types.append(t)
                p += 8 + n
            assert types[:2] == [1, 2] and set(types[2:]) <= {3} and got == body


@pytest.mark.gpu
@pytest.mark.parametrize("frame_max", [0, 9, 24, 64, 4096, 131072])
def test_deliveries_against_the_restated_render(frame_max):
    import torch

    from chanamq_b200 import decode as gd
    from chanamq_b200 import encode as ge

    rng = random.Random(100 + frame_max)
    stream, info = _publishes(rng, 300)
    data = np.frombuffer(stream, dtype=np.uint8)
    dec = gd.decode(data, np.array([0, len(data)], dtype=np.uint64))
    assert len(dec.recs) == len(info) and not dec.seg_err.any()
    # deliveries: every record 0..3 times, shuffled, with consumer tags of 0..255 bytes
    tags = [bytes(rng.randrange(33, 127) for _ in range(n)) for n in (0, 1, 5, 31, 32, 200, 255)]
    tag_off = np.cumsum([0] + [len(t) for t in tags])
    dels = [(r, rng.randrange(len(tags)), rng.getrandbits(64), rng.random() < 0.3, rng.randrange(1, 65535))
            for r in range(len(info)) for _ in range(rng.randrange(0, 4))]
    rng.shuffle(dels)
    darr = np.zeros(len(dels), dtype=ge.DELIVERY_DTYPE)
    for i, (r, t, dt, red, ch) in enumerate(dels):
        darr[i] = (dt, r, tag_off[t], ch, len(tags[t]), red, 0)
    want = [po.render_deliver(ch, tags[t], dt, red, info[r][0], info[r][1], info[r][2], info[r][3], frame_max) for r, t, dt, red, ch in dels]
    dev = torch.device("cuda:0")
    d_bytes = torch.from_numpy(data.copy()).to(dev)
    d_recs = torch.from_numpy(np.frombuffer(dec.recs.tobytes(), dtype=np.uint8).copy()).to(dev)
    d_tags = torch.from_numpy(np.frombuffer(b"".join(tags) + b"\0", dtype=np.uint8).copy()).to(dev)
    d_del = torch.from_numpy(np.frombuffer(darr.tobytes(), dtype=np.uint8).copy()).to(dev)
    total = sum(len(x) for x in want)
    d_out = torch.full((total + 16,), 0xAB, dtype=torch.uint8, device=dev)
    d_off = torch.zeros(len(dels) + 1, dtype=torch.int64, device=dev)
    d_st = torch.full((1,), 99, dtype=torch.int32, device=dev)
    ge.deliver_encode_device(d_bytes.data_ptr(), d_recs.data_ptr(), len(dec.recs), d_tags.data_ptr(), d_del.data_ptr(), len(dels),
                             frame_max, d_out.data_ptr(), total, d_off.data_ptr(), d_st.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert int(d_st[0]) == 0
    off = d_off.cpu().numpy()
    assert off.tolist() == np.cumsum([0] + [len(x) for x in want]).tolist()
    out = d_out.cpu().numpy().tobytes()
    for i, x in enumerate(want):
        assert out[off[i]:off[i + 1]] == x, (i, dels[i][0], frame_max)
    assert out[total:] == b"\xab" * 16   # nothing written past the end
    # the rendered stream is valid AMQP: the decoder assembles one command per delivery with the same body sizes
    back = gd.decode(np.frombuffer(out[:total], dtype=np.uint8), np.array([0, total], dtype=np.uint64))
    assert len(back.recs) == len(dels) and not back.seg_err.any()
    assert back.recs["body_size"].tolist() == [len(info[r][3]) for r, *_ in dels]


@pytest.mark.gpu
def test_encode_edge_cases():
    import torch

    from chanamq_b200 import decode as gd
    from chanamq_b200 import encode as ge

    stream = ab.message("ex", "a.b", b"x" * 50) + ab.other_method(60, 80, b"\x00" * 9)
    data = np.frombuffer(stream, dtype=np.uint8)
    dec = gd.decode(data, np.array([0, len(data)], dtype=np.uint64))
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d_bytes = torch.from_numpy(data.copy()).to(dev)
    d_recs = torch.from_numpy(np.frombuffer(dec.recs.tobytes(), dtype=np.uint8).copy()).to(dev)
    d_tags = torch.zeros(4, dtype=torch.uint8, device=dev)
    d_out = torch.full((4096,), 0xAB, dtype=torch.uint8, device=dev)
    d_st = torch.full((1,), 99, dtype=torch.int32, device=dev)

    def run(dels, cap):
        darr = np.zeros(max(len(dels), 1), dtype=ge.DELIVERY_DTYPE)
        for i, r in enumerate(dels):
            darr[i] = (1, r, 0, 1, 0, 0, 0)
        d_del = torch.from_numpy(np.frombuffer(darr.tobytes(), dtype=np.uint8).copy()).to(dev)
        d_off = torch.full((len(dels) + 1,), -1, dtype=torch.int64, device=dev)
        ge.deliver_encode_device(d_bytes.data_ptr(), d_recs.data_ptr(), len(dec.recs), d_tags.data_ptr(), d_del.data_ptr(), len(dels), 0,
                                 d_out.data_ptr(), cap, d_off.data_ptr(), d_st.data_ptr(), st)
        torch.cuda.synchronize()
        return int(d_st[0]), d_off.cpu().numpy().tolist()

    assert run([], 4096) == (0, [0])                                   # no deliveries
    one = len(po.render_deliver(1, b"", 1, False, b"ex", b"a.b", ab.content_header(50)[7:-1], b"x" * 50, 0))
    assert run([0, 0], 4096) == (0, [0, one, 2 * one])
    st_, off = run([0, 0], one)                                        # room for one only: the second is left out
    assert st_ == -3 and off == [0, one, 2 * one] and int(d_out[one]) in (0xAB, 1)
    assert run([1], 4096) == (-5, [0, 0])                              # record 1 is not a Basic.Publish
    assert run([7], 4096) == (-5, [0, 0])                              # no such record
