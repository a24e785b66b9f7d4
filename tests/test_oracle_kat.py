"""Pins the oracle against the reference's only known-answer test for routing:
TrieMatcher.main, chana-mq-server/.../engine/QueueMatcher.scala:75-139, plus the vectors SURVEY.md
A.1/A.8 froze from a line-faithful transliteration (Java split rules, the A.6 unbind defect)."""
import pytest

from oracle import pyoracle as po

S3 = "tmp_541b184c-6213-487c-a254-f1a8e3d461ab"
KAT_BINDS = [  # QM:87-94
    ("forex.*", "s0"), ("*.usd", "s0"), ("forex.eur", "s0"), ("*.eur", "s1"),
    ("forex.*", "s1"), ("trade", "s1"), ("*", "s2"), ("quote", S3),
]
KAT_EXPECT = {  # QM:105-114 (+ the unasserted println at QM:115-116)
    "forex.eur": ["s0", "s1"], "forex": ["s2"], "trade.jpy": [], "forex.jpy": ["s0", "s1"],
    "trade": ["s1", "s2"], "quote": ["s2", S3],
}


def _names(res):
    return [[n.decode() for n in r] if r is not None else None for r in res]


@pytest.mark.parametrize("mode", [po.COMPAT, po.SPEC])
@pytest.mark.parametrize("inplace", [False, True])
def test_triematcher_main_kat(mode, inplace):
    o = po.Oracle(po.TOPIC, mode, inplace=inplace)
    for k, q in KAT_BINDS:
        assert o.bind(k, q) == 0
    keys = list(KAT_EXPECT)
    assert _names(o.route(keys)) == [sorted(KAT_EXPECT[k]) for k in keys]
    for k, q in KAT_BINDS[:7]:  # QM:118-124: sub0..sub6
        assert o.unbind(k, q) == 0
    after = _names(o.route(keys))
    assert after[:5] == [[], [], [], [], []]  # QM:127-136
    assert after[5] == [S3]


def test_java_split_vectors():
    cases = {
        "": [b""], "a.": [b"a"], "a.b..": [b"a", b"b"], ".a": [b"", b"a"], "a..b": [b"a", b"", b"b"],
        ".": [], "..": [], "#": [b"#"], "a.#": [b"a", b"#"], "a": [b"a"],
    }
    for s, want in cases.items():
        got = po.split(s)
        if want == [b""]:
            assert po.lib().orc_split(b"", 0, None, 0, __import__("ctypes").byref(__import__("ctypes").c_uint32())) == 1
        else:
            assert got == want, s


def test_compat_edge_vectors():
    o = po.Oracle(po.TOPIC, po.COMPAT)
    for k, q in [("a.#", "qh"), ("*", "qs"), ("a.*.x", "qm"), ("a.*", "q2"), ("a", "qa")]:
        assert o.bind(k, q) == 0
    assert o.bind(".", "zz") == -1  # zero words: ArrayIndexOutOfBounds in the reference
    r = _names(o.route(["a.b", "a.#", "", "a..x", "a.", "a.b..", ".", "*"]))
    assert r[0] == ["q2"]            # '#' is a literal in the reference
    assert r[1] == ["q2", "qh"]
    assert r[2] == ["qs"]            # '*' matches the empty word
    assert r[3] == ["qm"]            # '*' matches an empty middle word; 'a.*' does not (3 words vs 2)
    assert r[4] == ["qa", "qs"]      # "a." == "a"
    assert r[5] == ["q2"]            # "a.b.." == "a.b"
    assert r[6] is None              # zero-word key
    assert r[7] == ["qs"]            # a key word "*" only rides the wildcard branch


def test_reference_unbind_defect_is_reproduced_by_the_oracle():
    o = po.Oracle(po.TOPIC, po.COMPAT)
    for k, q in [("a", "q1"), ("a.b", "q2"), ("*.x", "q3")]:
        o.bind(k, q)
    assert _names(o.route(["a.x", "a"])) == [["q3"], ["q1"]]
    o.unbind("a.b", "q2")
    assert _names(o.route(["a.x", "a"])) == [[], ["q1"]]  # q3 is lost: the reference's defect
    assert o.bind("a.c", "q4") == -2                      # and this subscribe never terminates there


def test_direct_and_fanout():  # QM:29-66
    d = po.Oracle(po.DIRECT)
    d.bind("a.b", "q1"); d.bind("a.*", "q2"); d.bind("a.b", "q3")
    assert _names(d.route(["a.b", "a.*", "a.c", "a.b."])) == [["q1", "q3"], ["q2"], [], []]
    f = po.Oracle(po.FANOUT)
    f.bind("x", "q1"); f.bind("y", "q2"); f.bind("x", "q1")
    assert _names(f.route(["", "whatever", "."])) == [["q1", "q2"]] * 3
    f.unbind("x", "q1")
    assert _names(f.route(["k"])) == [["q2"]]


def test_durable_flag_and_dedup():  # EX:300-302
    o = po.Oracle(po.TOPIC, po.COMPAT)
    o.bind("a.*", "q1", durable=False); o.bind("*.b", "q1", durable=False); o.bind("a.b", "q2", durable=True)
    blob, off = po.pack_keys(["a.b", "a.c", "z.z"])
    offs, ids, dur, err = o.route_packed(blob, off)
    assert offs.tolist() == [0, 2, 3, 3] and ids.tolist() == [0, 1, 0]
    assert dur.tolist() == [1, 0, 0] and err.tolist() == [0, 0, 0]
