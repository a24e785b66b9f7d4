"""Pins the decode oracle (oracle/amqp_decode.cpp, a restatement of FrameParser / CommandAssembler /
BasicProperties.readFrom / ValueReader) on the one framing vector the reference holds — the heartbeat frame
08 00 00 00 00 00 00 CE (chana-mq-base/.../model/Frame.scala:58-64) — and on hand-assembled frames for every
property and field type, the assembler's error cases and the chunk-carrying state of FrameParser.  The same
vectors are replayed through the CUDA path in tests/test_gpu_decode.py."""
import struct

import numpy as np
import pytest

from oracle import pyoracle as po
from tests import amqp_build as ab

F = dict(mandatory=1, immediate=2, persist=4, ttl=8, headers=16, props=32)


def dec(data, segs=None, chunk=0):
    data = np.frombuffer(bytes(data), dtype=np.uint8)
    segs = segs or [0, len(data)]
    return po.decode(data, np.array(segs, dtype=np.uint64), chunk)


def vectors():
    """name -> (bytes, expected) shared with the GPU test; expected = dict(n, err, consumed, checks=lambda recs,data)"""
    v = {}
    v["heartbeat_only"] = (bytes([8, 0, 0, 0, 0, 0, 0, 0xCE]), dict(n=0, err=0, consumed=8))
    rich_tbl = [("s", "S", b"str"), ("i", "I", -5), ("dec", "D", (2, 314)), ("ts", "T", 1700000000), ("nest", "F", [("x", "t", True)]),
                ("arr", "A", [("I", 1), ("S", b"zz"), ("F", [("k", "V", None)])]), ("b", "b", -1), ("dbl", "d", 1.5), ("flt", "f", 2.5),
                ("lng", "l", 1 << 40), ("sh", "s", -2), ("bool", "t", False), ("bytes", "x", b"\x00\x01"), ("void", "V", None),
                ("s", "S", b"dup")]
    m1 = ab.message("ex", "a.b.c", b"B" * 10, channel=3, contentType="text/plain", contentEncoding="utf8", headers=rich_tbl,
                    deliveryMode=2, priority=4, correlationId="cid", replyTo="rq", expiration="60000", messageId="mid",
                    timestamp=1700000001, type="t", userId="guest", appId="app", clusterId="cl")
    v["all_properties_all_field_types"] = (m1, dict(n=1, err=0, consumed=len(m1), flags=F["persist"] | F["ttl"] | F["headers"] | F["props"],
                                                    ttl=60000, key=b"a.b.c", exch=b"ex", body_size=10, channel=3))
    m2 = ab.publish_method("e", "k", True, True) + ab.content_header(0, deliveryMode=1)
    v["zero_body_completes_at_header_mandatory_immediate"] = (m2, dict(n=1, err=0, consumed=len(m2), body_size=0,
                                                                       flags=F["mandatory"] | F["immediate"] | F["props"]))
    m3 = (ab.publish_method("e", "k") + ab.heartbeat() + ab.content_header(7) + ab.body_frame(b"") + ab.body_frame(b"abc") +
          ab.heartbeat() + ab.body_frame(b"defg"))
    v["multi_frame_body_with_heartbeats_and_empty_body_frames"] = (m3, dict(n=1, err=0, consumed=len(m3), body_size=7))
    m4 = ab.other_method(20, 10, shortstr_pad()) + ab.message("e", "k2", b"x") + ab.other_method(60, 80, struct.pack(">QB", 9, 0))
    v["other_methods_are_passed_through"] = (m4, dict(n=3, err=0, consumed=len(m4), kinds=[0, 1, 0]))
    whole = ab.message("e", "k", b"12345")
    v["partial_tail_is_left_unconsumed"] = (whole + whole[:-3], dict(n=1, err=0, consumed=len(whole)))
    v["partial_after_header"] = (whole + ab.publish_method("e", "k") + ab.content_header(5), dict(n=1, err=0, consumed=len(whole)))
    for name, ttl_s, ok, val in [("ttl_plus", "+15", True, 15), ("ttl_neg", "-3", True, -3), ("ttl_junk", "12x", False, 0),
                                 ("ttl_empty", "", False, 0), ("ttl_space", " 5", False, 0), ("ttl_max", "9223372036854775807", True, (1 << 63) - 1),
                                 ("ttl_overflow", "9223372036854775808", False, 0), ("ttl_min", "-9223372036854775808", True, -(1 << 63))]:
        m = ab.message("e", "k", b"", expiration=ttl_s)
        v[name] = (m, dict(n=1, err=0, consumed=len(m), flags=F["props"] | (F["ttl"] if ok else 0), ttl=val))
    # --- errors: the segment stops at the first one; earlier commands are still reported -------------------
    bad_end = whole + ab.frame(ab.METHOD, 1, struct.pack(">HHH", 60, 40, 0) + ab.shortstr("e") + ab.shortstr("k") + b"\0", end=0xCD)
    v["bad_end_marker_is_frame_error_501"] = (bad_end, dict(n=1, err=501, consumed=len(whole), err_off=len(whole)))
    v["header_without_method"] = (ab.content_header(1), dict(n=0, err=501, consumed=0, err_off=0))
    v["method_where_header_expected"] = (ab.publish_method("e", "k") + ab.publish_method("e", "k"), dict(n=0, err=501, consumed=0))
    v["body_overshoot"] = (ab.publish_method("e", "k") + ab.content_header(3) + ab.body_frame(b"abcd"), dict(n=0, err=501, consumed=0))
    v["unknown_class_in_method_throws"] = (whole + ab.other_method(77, 1), dict(n=1, err=1, consumed=len(whole), err_off=len(whole)))
    v["unknown_class_in_header_throws"] = (ab.publish_method("e", "k") + ab.content_header(1, class_id=61), dict(n=0, err=1, consumed=0))
    v["continuation_flag_throws"] = (ab.publish_method("e", "k") + ab.content_header(1, flags_extra=1), dict(n=0, err=1, consumed=0))
    trunc = ab.frame(ab.METHOD, 1, struct.pack(">HHH", 60, 40, 0) + ab.shortstr("e") + bytes([200]) + b"short")
    v["truncated_shortstr_throws"] = (trunc, dict(n=0, err=1, consumed=0))
    v["bad_table_type_throws"] = (ab.publish_method("e", "k") + ab.content_header(0, headers=struct.pack(">I", 4) + ab.shortstr("n") + b"?" + b"\0"),
                                  dict(n=0, err=1, consumed=0))
    v["negative_frame_size_throws"] = (whole + struct.pack(">BHI", 1, 1, 0x80000001) + b"\0" * 16, dict(n=1, err=1, consumed=len(whole)))
    over = ab.publish_method("e", "k") + ab.content_header(0, headers=struct.pack(">I", 500) + ab.shortstr("n") + ab.field("I", 7))
    v["table_length_overrunning_the_payload_is_clamped"] = (over, dict(n=1, err=0, consumed=len(over), flags=F["headers"] | F["props"]))
    return v


def shortstr_pad():
    return ab.shortstr("") + b"\0" * 4


def check_vector(name, data, want, recs, consumed, err, err_off):
    assert len(recs) == want["n"], name
    assert int(err[0]) == want["err"], name
    assert int(consumed[0]) == want["consumed"], name
    if "err_off" in want:
        assert int(err_off[0]) == want["err_off"], name
    if want["n"] and "kinds" not in want:
        r = recs[0]
        if "flags" in want:
            assert int(r["flags"]) == want["flags"], (name, int(r["flags"]))
        if "ttl" in want:
            assert int(r["ttl"]) == want["ttl"], name
        if "key" in want:
            assert bytes(data[r["key_off"]:r["key_off"] + r["key_len"]]) == want["key"], name
            assert bytes(data[r["exch_off"]:r["exch_off"] + r["exch_len"]]) == want["exch"], name
        if "body_size" in want:
            assert int(r["body_size"]) == want["body_size"], name
        if "channel" in want:
            assert int(r["channel"]) == want["channel"], name
    if "kinds" in want:
        assert recs["kind"].tolist() == want["kinds"], name


@pytest.mark.parametrize("chunk", [0, 1, 5, 64])
def test_vectors(chunk):
    for name, (data, want) in vectors().items():
        recs, consumed, err, err_off = dec(data, chunk=chunk)
        check_vector(name, data, want, recs, consumed, err, err_off)


def test_record_details():
    data, _ = vectors()["all_properties_all_field_types"]
    recs, *_ = dec(data)
    r = recs[0]
    tbl = data[r["hdr_tbl_off"]:r["hdr_tbl_off"] + r["hdr_tbl_len"]]
    assert tbl.startswith(ab.shortstr("s") + b"S") and tbl.endswith(ab.shortstr("s") + ab.field("S", b"dup"))
    assert data[r["body_off"]:r["body_off"] + 10] == b"B" * 10 and r["method_off"] == 0 and r["seg"] == 0
    data, _ = vectors()["other_methods_are_passed_through"]
    recs, *_ = dec(data)
    assert (int(recs[0]["key_off"]), int(recs[0]["key_len"])) == (20, 10) and (int(recs[2]["key_off"]), int(recs[2]["key_len"])) == (60, 80)
    assert data[recs[2]["exch_off"]:recs[2]["exch_off"] + recs[2]["exch_len"]] == struct.pack(">HHQB", 60, 80, 9, 0)
    data, _ = vectors()["multi_frame_body_with_heartbeats_and_empty_body_frames"]
    recs, *_ = dec(data)
    assert data[recs[0]["body_off"]:recs[0]["body_off"] + 3] == b"abc"


def test_segments_are_independent():
    a, b = ab.message("e", "one", b"1"), ab.message("e", "two", b"22") + ab.heartbeat()
    bad = ab.content_header(1)
    data = a + bad + b
    recs, consumed, err, err_off = dec(data, [0, len(a), len(a) + len(bad), len(data)])
    assert len(recs) == 2 and recs["seg"].tolist() == [0, 2]
    assert consumed.tolist() == [len(a), 0, len(b)] and err.tolist() == [0, 501, 0] and int(err_off[1]) == len(a)


def test_generated_stream_roundtrip():
    from chanamq_b200.workload import TopicWorkload, amqp_stream

    w = TopicWorkload(5)
    data, seg, n_msgs = amqp_stream(w, 2_000_000, 16, body_bytes=512, stream=1)
    recs, consumed, err, err_off = po.decode(data, seg)
    assert n_msgs > 3000 and len(recs) == n_msgs and not err.any()
    assert np.array_equal(consumed, np.diff(seg)) and (recs["kind"] == 1).all()
    assert (recs["body_size"] == 512).all() and ((recs["flags"] & 32) != 0).all()
    rich = (recs["flags"] & 16) != 0
    assert 0.05 < rich.mean() < 0.15 and ((recs["flags"][rich] & 8) != 0).all()
    # chunked feeding (TCP reads of 1500 bytes) gives the same records: FrameParser carries its tail
    recs2, consumed2, err2, _ = po.decode(data, seg, chunk=1500)
    assert recs2.tobytes() == recs.tobytes() and np.array_equal(consumed2, consumed)
    # the routing keys are the generator's keys
    keys = [bytes(data[r["key_off"]:r["key_off"] + r["key_len"]]) for r in recs[:50]]
    assert all(3 <= len(k.split(b".")) <= 6 for k in keys)
