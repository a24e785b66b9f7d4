"""GPU parity tests proper: every case goes through the C ABI (libchanagpu.so) and is compared bit
for bit with the CPU oracle (oracle/), which restates the reference's TrieMatcher / Direct /
Fanout matchers and the dedup of ExchangeEntity.scala:297-302.  Integer work: the bar is equality."""
import random

import numpy as np
import pytest

from chanamq_b200 import CHQ_DIRECT, CHQ_FANOUT, CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC, ChqError, GpuMatcher
from chanamq_b200.workload import TopicWorkload, queue_names_packed
from oracle import pyoracle as po
from tests.helpers import assert_same_routing, check_csr_invariants, load_both, pack
from tests.test_oracle_kat import KAT_BINDS, KAT_EXPECT, S3

pytestmark = pytest.mark.gpu


def _names(res):
    return [[n.decode() for n in r] if r is not None else None for r in res]


@pytest.mark.parametrize("mode", [CHQ_MODE_COMPAT, CHQ_MODE_SPEC])
def test_triematcher_main_kat(mode):  # QueueMatcher.scala:75-139 through the C ABI
    m = GpuMatcher(CHQ_TOPIC, mode)
    subs = [m.bind(k, q) for k, q in KAT_BINDS]
    keys = list(KAT_EXPECT)
    assert _names(m.lookup_all(keys)) == [sorted(KAT_EXPECT[k]) for k in keys]
    assert m.lookup("forex.eur") == {b"s0", b"s1"}
    for s in subs[:7]:
        m.unsubscribe(s)
    after = _names(m.lookup_all(keys))
    assert after[:5] == [[], [], [], [], []] and after[5] == [S3]


def test_compat_edge_vectors():
    m, o = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT), po.Oracle(po.TOPIC, po.COMPAT)
    load_both(m, o, [("a.#", "qh", 0), ("*", "qs", 0), ("a.*.x", "qm", 1), ("a.*", "q2", 0), ("a", "qa", 0),
                     (".", "zz", 0), ("..", "zz", 0), ("", "qe", 1), ("a..b", "qd", 0), (".a", "ql", 0)])
    assert m.try_bind(".", "zz") == -1
    keys = ["a.b", "a.#", "", "a..x", "a.", "a.b..", ".", "*", "..", "a..b", ".a", "a.*", "x" * 255, "a." * 127 + "a"]
    blob, off = pack(keys)
    r = assert_same_routing(m, o, blob, off)
    assert r.err.tolist()[6] == 1 and r.err.tolist()[8] == 1
    with pytest.raises(IndexError):
        m.lookup(".")


def test_key_longer_than_shortstr_is_flagged():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    m.bind("*", "q")
    assert m.try_bind("x" * 256, "q") == -2
    blob, off = pack(["x" * 256, "x"])
    r = m.route_packed(blob, off)
    assert r.err.tolist() == [2, 0] and r.offsets.tolist() == [0, 0, 1]


def test_unbind_is_clean_set_removal_not_the_reference_defect():  # deliberately NOT reproduced
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    for k, q in [("a", "q1"), ("a.b", "q2"), ("*.x", "q3")]:
        m.bind(k, q)
    assert _names(m.lookup_all(["a.x", "a"])) == [["q3"], ["q1"]]
    m.unbind("a.b", "q2")
    assert _names(m.lookup_all(["a.x", "a"])) == [["q3"], ["q1"]]  # the reference would now answer [] for a.x
    m.bind("a.c", "q4")
    assert _names(m.lookup_all(["a.c"])) == [["q4"]]
    m.unbind_queue("q3")
    assert _names(m.lookup_all(["a.x"])) == [[]]


def test_empty_table_and_empty_batch():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT)
    r = m.route_packed(*pack(["a", "b.c"]))
    assert r.offsets.tolist() == [0, 0, 0] and len(r.queue_ids) == 0
    m.bind("a", "q")
    r = m.route_packed(np.zeros(1, np.uint8), np.zeros(1, np.uint32))
    assert r.offsets.tolist() == [0]


def _rand_topic(rng, alphabet, maxw, wild, pw=0.35):
    n = rng.randint(1, maxw)
    return ".".join(rng.choice(wild) if rng.random() < pw else rng.choice(alphabet) for _ in range(n))


@pytest.mark.parametrize("mode,limits", [(CHQ_MODE_COMPAT, (0, 0)), (CHQ_MODE_SPEC, (0, 0)), (CHQ_MODE_SPEC, (1, 0)),
                                         (CHQ_MODE_SPEC, (2, 3)), (CHQ_MODE_COMPAT, (3, 1)), (CHQ_MODE_SPEC, (8, 5))])
def test_random_small_tables(mode, limits):
    """dense little tables: many shapes, shared queues (dedup), '#' anywhere, empty words; the limits
    force the shared-memory merge and the global-scratch slow path"""
    rng = random.Random(1000 + mode * 10 + limits[0])
    alpha = ["a", "b", "c", "", "dd", "eeeeeeeeee", "#x"]
    omode = po.SPEC if mode == CHQ_MODE_SPEC else po.COMPAT
    for trial in range(12):
        m, o = GpuMatcher(CHQ_TOPIC, mode), po.Oracle(po.TOPIC, omode)
        m.debug_limits(*limits)
        nq = rng.choice([3, 17, 60])
        binds = [(_rand_topic(rng, alpha, 6, ["*", "#", "#", "*"]), "q%03d" % rng.randrange(nq), rng.random() < 0.3)
                 for _ in range(rng.choice([5, 80, 400]))]
        load_both(m, o, binds)
        keys = [_rand_topic(rng, alpha + ["*", "#"], 7, ["zz"], 0.1) for _ in range(700)]
        blob, off = pack(keys)
        r = assert_same_routing(m, o, blob, off, what="trial %d" % trial)
        check_csr_invariants(r, m.n_queues())


@pytest.mark.parametrize("mode", [CHQ_MODE_COMPAT, CHQ_MODE_SPEC])
def test_runs_of_dots_make_up_to_255_words(mode):
    """a shortstr of 255 bytes holds up to 255 words when most of them are empty ("a....b"): patterns whose
    remainder behind the 16 indexed words is ~200 items long, keys of 129..255 words"""
    m, o = GpuMatcher(CHQ_TOPIC, mode), po.Oracle(po.TOPIC, po.SPEC if mode == CHQ_MODE_SPEC else po.COMPAT)
    binds = [("#" + "." * 200 + "a", "q1", False), ("a" + "." * 128 + "b", "q2", True), ("a" + "." * 253 + "b", "q3", False),
             ("a.#." + "." * 150 + "b", "q4", False), ("*" + "." * 128 + "*", "q5", False), ("a.#", "q6", False),
             ("#.b", "q7", False), ("a" + "." * 100 + "#" + "." * 100 + "b", "q8", True)]
    load_both(m, o, binds)
    keys = ["a" + "." * 128 + "b", "a" + "." * 253 + "b", "x" + "." * 200 + "a", "." * 200 + "a", "a" + "." * 152 + "b",
            "a" + "." * 201 + "b", "a" + "." * 127 + "b", "a" + "." * 129 + "b", "a.b", "a" + "." * 254, "." * 254 + "b"]
    r = assert_same_routing(m, o, *pack(keys))
    check_csr_invariants(r, m.n_queues())


def test_long_patterns_and_long_keys_spec():
    """patterns beyond the 16 indexed words, '#' in the middle of long keys (word stride 128)"""
    rng = random.Random(5)
    m, o = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC), po.Oracle(po.TOPIC, po.SPEC)
    words = ["a", "b", "c"]
    binds = []
    for i in range(120):
        n = rng.randint(1, 40)
        binds.append((".".join(rng.choice(words + ["*", "#"]) for _ in range(n)), "q%02d" % (i % 30), False))
    binds += [("a.#.b", "mid", True), ("#.c", "tail", False), ("#", "all", False), ("a.#.b.#.c", "two", False),
              (".".join(["a"] * 20), "long", False), (".".join(["a"] * 17 + ["#"]), "longh", False)]
    load_both(m, o, binds)
    keys = [".".join(rng.choice(words) for _ in range(rng.randint(1, 60))) for _ in range(1500)]
    keys += [".".join(["a"] * 20), ".".join(["a"] * 19), ".".join(["a"] * 128), "a.x.y.z.w.b"]
    r = assert_same_routing(m, o, *pack(keys))
    check_csr_invariants(r, m.n_queues())


def test_direct_and_fanout_exchanges():  # QueueMatcher.scala:29-66
    d, od = GpuMatcher(CHQ_DIRECT), po.Oracle(po.DIRECT)
    load_both(d, od, [("a.b", "q1", 0), ("a.*", "q2", 1), ("a.b", "q3", 0), (".", "q4", 0), ("", "q5", 0)])
    assert_same_routing(d, od, *pack(["a.b", "a.*", "a.c", "a.b.", ".", "", "a"]))
    f, of = GpuMatcher(CHQ_FANOUT), po.Oracle(po.FANOUT)
    load_both(f, of, [("x", "q1", 0), ("y", "q2", 1), ("x", "q1", 0)])
    assert_same_routing(f, of, *pack(["", "whatever", ".", "a.b"]))
    f.unbind("y", "q2", True); of.unbind("y", "q2", True)
    assert_same_routing(f, of, *pack(["k"]))


def _load_workload(seed, p, n_queues, mode, hash_tail):
    w = TopicWorkload(seed)
    pb, poff, q, d = w.patterns(p, n_queues, hash_tail)
    qb, qoff = queue_names_packed(q)
    m = GpuMatcher(CHQ_TOPIC, mode)
    o = po.Oracle(po.TOPIC, po.SPEC if mode == CHQ_MODE_SPEC else po.COMPAT, inplace=True)
    m.bind_many(pb, poff, qb, qoff, d)
    assert o.bind_many(pb, poff, qb, qoff, d) == 0
    m.commit()
    return w, m, o


def test_config1_topic_1k_bindings_100k_publishes():
    """BASELINE.json configs[0]: topic, 1 k bindings, 100 k publishes — the whole case against the oracle"""
    w, m, o = _load_workload(1, 1000, 1000, CHQ_MODE_COMPAT, False)
    blob, off = w.keys(100_000, stream=0)
    r = assert_same_routing(m, o, blob, off, threads=4)
    check_csr_invariants(r, m.n_queues())


@pytest.mark.parametrize("mode,hash_tail,nq", [(CHQ_MODE_COMPAT, True, 100_000), (CHQ_MODE_SPEC, True, 100_000),
                                               (CHQ_MODE_COMPAT, False, 25_000), (CHQ_MODE_SPEC, True, 25_000)])
def test_config2_topic_100k_bindings(mode, hash_tail, nq):
    """configs[1] shape (100 k bindings, mixed */#) on 300 k keys; nq = P/4 is the dedup variant.
    In COMPAT mode the '#' of the generated patterns is an ordinary word, as in the reference."""
    w, m, o = _load_workload(2, 100_000, nq, mode, hash_tail)
    for stream in range(2):
        blob, off = w.keys(150_000, stream=stream)
        r = assert_same_routing(m, o, blob, off, threads=4, what="stream %d" % stream)
        check_csr_invariants(r, m.n_queues())
    st = m.stats()
    assert st["bindings"] > 99_000 and st["wstride"] == 8


@pytest.mark.parametrize("mode", [CHQ_MODE_COMPAT, CHQ_MODE_SPEC])
def test_full_size_1m_bindings_properties_and_sample(mode):
    """configs[3] table (1 M bindings) at one full batch of 2^20 keys, in the benchmark's own mode (COMPAT) and with
    the '#' wildcard: CSR invariants over the whole result, idempotence, and exact equality with the oracle on a
    120 k-key sample"""
    w, m, o = _load_workload(4, 1_000_000, 1_000_000, mode, True)
    blob, off = w.keys(1 << 20, stream=3)
    r1 = m.route_packed(blob, off)
    check_csr_invariants(r1, m.n_queues())
    r2 = m.route_packed(blob, off)
    assert all(np.array_equal(a, b) for a, b in zip(r1, r2))
    n_s = 120_000
    sblob, soff = blob[: off[n_s]], off[: n_s + 1]
    rs = assert_same_routing(m, o, sblob, soff, threads=8)
    assert np.array_equal(rs.offsets, r1.offsets[: n_s + 1]) and np.array_equal(rs.queue_ids, r1.queue_ids[: rs.offsets[-1]])
    # linearity: routing the two halves separately concatenates to the whole
    half = 1 << 19
    ra = m.route_packed(blob[: off[half]], off[: half + 1])
    rb = m.route_packed(blob[off[half]:], off[half:] - off[half])
    assert np.array_equal(np.concatenate([ra.queue_ids, rb.queue_ids]), r1.queue_ids)
    assert np.array_equal(np.concatenate([ra.offsets[:-1], rb.offsets + ra.offsets[-1]]), r1.offsets)


def test_pipelined_slots_match_the_blocking_call():
    w, m, o = _load_workload(2, 20_000, 5_000, CHQ_MODE_SPEC, True)
    batches = [w.keys(30_000 + 17 * i, stream=10 + i) for i in range(6)]
    want = [m.route_packed(b, f) for b, f in batches]
    got = {}
    for i, (b, f) in enumerate(batches):
        slot = i % 3
        if i >= 3:
            got[i - 3] = m.wait(slot, copy=True)
        m.submit(slot, b.ctypes.data, f.ctypes.data, len(f) - 1)
    for i in range(3, 6):
        got[i] = m.wait(i % 3, copy=True)
    for i in range(6):
        assert all(np.array_equal(a, b) for a, b in zip(want[i], got[i])), i


def test_ids_as_three_bytes_equal_the_uint32_result():  # chq_set_ids_format(CHQ_IDS_U24), pack.cu
    import torch
    from chanamq_b200._lib import CHQ_IDS_U24, CHQ_IDS_U32
    w, m, o = _load_workload(2, 100_000, 100_000, CHQ_MODE_COMPAT, True)  # queue ids up to 99 999: all three bytes in use
    for n_keys in (0, 1, 7, 5_000, 60_001):
        blob, off = w.keys(n_keys, stream=5)
        m.set_ids_format(CHQ_IDS_U32)
        want = assert_same_routing(m, o, blob, off, threads=4) if n_keys else m.route_packed(blob, off)
        m.set_ids_format(CHQ_IDS_U24)
        for _ in range(2):  # twice: the second batch reuses the slot's buffers
            got = m.route_packed(blob, off)
            assert got.queue_ids.dtype == np.uint8 and len(got.queue_ids) == 3 * len(want.queue_ids)
            assert np.array_equal(got.offsets, want.offsets) and np.array_equal(got.err, want.err)
            assert np.array_equal(got.any_durable, want.any_durable)
            assert np.array_equal(m.unpack_u24(got.queue_ids), want.queue_ids)
        # pipelined slots keep the format they were submitted with
        hb = torch.from_numpy(np.concatenate([blob, np.zeros(1, np.uint8)])).pin_memory()
        ho = torch.from_numpy(off.view(np.int32).copy()).pin_memory()
        m.submit(0, hb.data_ptr(), ho.data_ptr(), n_keys)
        m.set_ids_format(CHQ_IDS_U32)
        m.submit(1, hb.data_ptr(), ho.data_ptr(), n_keys)
        r0, r1 = m.wait(0, copy=True), m.wait(1, copy=True)
        assert r0.queue_ids.dtype == np.uint8 and np.array_equal(m.unpack_u24(r0.queue_ids), want.queue_ids)
        assert r1.queue_ids.dtype == np.uint32 and np.array_equal(r1.queue_ids, want.queue_ids)
    assert int(want.queue_ids.max()) >= 1 << 16
    m.close()


def test_device_entry_point_with_torch_buffers():
    torch = pytest.importorskip("torch")
    w, m, o = _load_workload(2, 50_000, 50_000, CHQ_MODE_SPEC, True)
    blob, off = w.keys(100_000, stream=5)
    n = len(off) - 1
    dev = torch.device("cuda:0")
    d_blob = torch.from_numpy(blob.copy()).to(dev)
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    d_offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
    cap = 4_000_000
    d_ids = torch.empty(cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    d_err = torch.empty(n, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        m.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), cap,
                       d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
    torch.cuda.synchronize()
    total, status = int(d_tot[0].item()), int(d_tot[1].item()) & 0xFFFFFFFF
    assert status == 0
    o_off, o_ids, o_dur, o_err = o.route_packed(blob, off, 4)
    assert total == len(o_ids)
    assert np.array_equal(d_offsets.cpu().numpy().view(np.uint32), o_off)
    assert np.array_equal(d_ids[:total].cpu().numpy().view(np.uint32), o_ids)
    assert np.array_equal(d_dur.cpu().numpy(), o_dur)
    # too small an id buffer: flagged, nothing written past it
    d_ids2 = torch.full((1001,), -7, dtype=torch.int32, device=dev)
    m.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, d_offsets.data_ptr(), d_ids2.data_ptr(), 1000,
                   d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
    torch.cuda.synchronize()
    assert (int(d_tot[1].item()) & 0xFFFFFFFF) == (-3 & 0xFFFFFFFF) and int(d_ids2[1000].item()) == -7
    fit = int(o_off[o_off <= 1000].max())  # messages that fit entirely are still written
    assert fit > 0 and np.array_equal(d_ids2[:fit].cpu().numpy().view(np.uint32), o_ids[:fit])


@pytest.mark.parametrize("n", [1, 31, 32, 127, 128, 129, 255, 257, 1000])
def test_batch_sizes_around_tile_edges(n):
    w, m, o = _load_workload(2, 20_000, 5_000, CHQ_MODE_SPEC, True)
    blob, off = w.keys(n, stream=100 + n)
    r = assert_same_routing(m, o, blob, off)
    check_csr_invariants(r, m.n_queues())


def test_slow_path_arena_grows_and_retries(monkeypatch):
    """host-level calls re-run a batch whose slow path ran out of scratch (CHQ_E_SCRATCH) with a larger arena"""
    monkeypatch.setenv("CHQ_SCRATCH_IDS", "64")
    w, m, o = _load_workload(2, 20_000, 2_000, CHQ_MODE_SPEC, True)
    m.debug_limits(1, 1)  # every multi-list message takes the global-scratch path
    blob, off = w.keys(40_000, stream=7)
    r = assert_same_routing(m, o, blob, off, threads=4)
    assert len(r.queue_ids) > 10_000


@pytest.mark.parametrize("mode", [CHQ_MODE_COMPAT, CHQ_MODE_SPEC])
def test_word_lengths_around_the_8_byte_chunks(mode):
    """the tokeniser reads keys as aligned 32-bit words and hashes eight bytes at a time: words of 0..40
    bytes, words that only differ after byte 8 / 16, NUL and high bytes, every key alignment"""
    rng = random.Random(77 + mode)
    stems = [b"", b"a", b"abc", b"abcdefg", b"abcdefgh", b"abcdefghi", b"abcdefgh\x00", b"abcdefghijklmnop",
             b"abcdefghijklmnopq", b"abcdefghijklmnoX", b"abcdefghijklmnopqrstuvwxyz0123456789-_", b"\x00", b"\x00\x00",
             b"\xff\xfe", b"a\x00b", b"..".replace(b".", b"-"), b"#", b"*x"]
    def topic(wild):
        n = rng.randint(1, 6)
        return b".".join(rng.choice(wild) if rng.random() < 0.25 else rng.choice(stems) for _ in range(n))
    m, o = GpuMatcher(CHQ_TOPIC, mode), po.Oracle(po.TOPIC, po.SPEC if mode == CHQ_MODE_SPEC else po.COMPAT)
    load_both(m, o, [(topic([b"*", b"#"]), "q%02d" % rng.randrange(40), rng.random() < 0.5) for _ in range(600)])
    keys = [topic([b"zzzzzzzzzzzz"]) for _ in range(3000)]
    r = assert_same_routing(m, o, *pack(keys))
    assert len(r.queue_ids) > 1000
    d, od = GpuMatcher(CHQ_DIRECT), po.Oracle(po.DIRECT)   # whole key = one word, up to 255 bytes
    long_keys = [bytes(rng.choice(b"ab.") for _ in range(rng.choice([0, 1, 7, 8, 9, 15, 16, 17, 64, 200, 255]))) for _ in range(300)]
    load_both(d, od, [(k, "q%d" % (i % 7), False) for i, k in enumerate(long_keys[:150])])
    assert_same_routing(d, od, *pack(long_keys + [k + b"x" for k in long_keys[:50] if len(k) < 255]))


def test_misaligned_key_blob_on_the_device():
    """chq_route_device with the key bytes starting at every address modulo 4"""
    torch = pytest.importorskip("torch")
    w, m, o = _load_workload(2, 20_000, 20_000, CHQ_MODE_COMPAT, True)
    blob, off = w.keys(5_000, stream=9)
    n = len(off) - 1
    o_off, o_ids, o_dur, o_err = o.route_packed(blob, off, 2)
    dev = torch.device("cuda:0")
    d_off = torch.from_numpy(off.view(np.int32).copy()).to(dev)
    d_offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_ids = torch.empty(len(o_ids) + 1, dtype=torch.int32, device=dev)
    d_dur = torch.empty(n, dtype=torch.uint8, device=dev)
    d_err = torch.empty(n, dtype=torch.uint8, device=dev)
    d_tot = torch.zeros(2, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for shift in range(4):
        buf = torch.zeros(len(blob) + 8, dtype=torch.uint8, device=dev)
        buf[shift:shift + len(blob)] = torch.from_numpy(blob.copy()).to(dev)
        buf[shift + len(blob):] = ord(".")   # whatever follows the blob must not leak into the last key
        m.route_device(buf.data_ptr() + shift, d_off.data_ptr(), n, d_offsets.data_ptr(), d_ids.data_ptr(), len(o_ids) + 1,
                       d_dur.data_ptr(), d_err.data_ptr(), d_tot.data_ptr(), d_tot.data_ptr() + 8, st)
        torch.cuda.synchronize()
        assert (int(d_tot[1].item()) & 0xFFFFFFFF) == 0 and int(d_tot[0].item()) == len(o_ids)
        assert np.array_equal(d_offsets.cpu().numpy().view(np.uint32), o_off), shift
        assert np.array_equal(d_ids[:len(o_ids)].cpu().numpy().view(np.uint32), o_ids), shift


@pytest.mark.parametrize("max_small", [0, 1, 40])
def test_several_long_lists_without_shared_queues(max_small):
    """messages whose matched lists are all long and disjoint: merged in place by the multi-way path
    (list lengths around the 32-id window, up to 8 lists)"""
    rng = random.Random(31 + max_small)
    m, o = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC), po.Oracle(po.TOPIC, po.SPEC)
    m.debug_limits(0, max_small)
    pats = ["a.#", "a.b.#", "a.*.c", "a.b.c", "*.b.c", "a.b.*", "#", "*.*.c", "a.b.c.#", "x.y", "x.*", "x.#"]
    sizes = [500, 333, 64, 32, 31, 33, 129, 1, 97, 1000, 999, 65]
    binds, q = [], 0
    order = list(range(200000))
    rng.shuffle(order)   # queue names in random order, so the lists interleave
    for pat, sz in zip(pats, sizes):
        for _ in range(sz):
            binds.append((pat, "q%06d" % order[q], rng.random() < 0.2))
            q += 1
    rng.shuffle(binds)
    load_both(m, o, binds)
    keys = ["a.b.c", "a.b", "a", "a.z.c", "x.y", "x.z", "x", "q", "a.b.c.d", "z.b.c", "a.b.d"] * 40
    r = assert_same_routing(m, o, *pack(keys))
    check_csr_invariants(r, m.n_queues())
    cnt = np.diff(r.offsets.astype(np.int64))
    assert int(cnt[0]) == sum(sizes[:9]) and int(cnt[4]) == sizes[9] + sizes[10] + sizes[11] + sizes[6]
