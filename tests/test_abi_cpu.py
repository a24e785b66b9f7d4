"""CPU-side checks of the product library (no compute calls: this container has no GPU): the C-ABI
library loads and exports every symbol include/chanamq_gpu.h declares, the host-side table compiler
behaves (host-only tables, device = -1), and the product refuses to route without CUDA — there is
no CPU fallback."""
import ctypes as C

import numpy as np
import pytest

from chanamq_b200 import CHQ_DIRECT, CHQ_FANOUT, CHQ_MODE_COMPAT, CHQ_MODE_SPEC, CHQ_TOPIC, ChqError, GpuMatcher
from chanamq_b200 import _lib
from chanamq_b200.workload import TopicWorkload, queue_names_packed
from tests.test_oracle_kat import KAT_BINDS


def test_library_exports_every_declared_symbol():
    L = _lib.load_library()
    declared = _lib.declared_symbols()
    assert len(declared) >= 18
    missing = [s for s in declared if not hasattr(L, s)]
    assert missing == []


def test_no_cpu_fallback():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT, device=-1)
    m.bind("a.*", "q")
    m.commit()
    with pytest.raises(ChqError) as e:
        m.lookup_all(["a.b"])
    assert e.value.code == _lib.CHQ_E_CUDA


def test_host_compile_of_the_reference_kat_table():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT, device=-1)
    for k, q in KAT_BINDS:
        m.bind(k, q)
    m.bind("forex.*", "s0")  # identical pair again: no-op (QueueMatcher.scala:443-446)
    st = m.stats()
    assert st["bindings"] == 8 and st["queues"] == 4 and st["groups"] == 7 and st["shapes"] == 5
    assert m.queue_names() == [b"s0", b"s1", b"s2", b"tmp_541b184c-6213-487c-a254-f1a8e3d461ab"]
    assert m.try_bind(".", "x") == _lib.CHQ_E_ZERO_WORDS and m.try_bind("..", "x") == _lib.CHQ_E_ZERO_WORDS
    assert m.try_bind("k" * 256, "x") == _lib.CHQ_E_KEY_TOO_LONG
    m.unbind_queue("s0")
    # applied in place: the bindings are gone, the queue keeps its id (and its place among the names) ...
    assert m.stats()["bindings"] == 5 and m.stats()["incremental_commits"] == 1 and b"s0" in m.queue_names()
    m.commit_full()   # ... until the whole set is compiled again
    assert m.stats()["bindings"] == 5 and b"s0" not in m.queue_names()


def test_shapes_and_word_stride():
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC, device=-1)
    for k in ["a.b.c", "a.*.c", "*.b.c", "a.b.#", "a.#", "x.y.z.w.v.u"]:
        m.bind(k, "q")
    st = m.stats()
    assert st["shapes"] == 6 and st["wstride"] == 8
    m.bind("a.#.b", "q")           # '#' in the middle: remainder verified by the NFA, all words kept
    assert m.stats()["wstride"] == 256   # 255 bytes hold up to 255 words
    c = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT, device=-1)
    c.bind("a.#.b", "q")           # COMPAT: '#' is a word (QueueMatcher.scala:70)
    assert c.stats()["wstride"] == 4 and c.stats()["shapes"] == 1
    d = GpuMatcher(CHQ_DIRECT, device=-1)
    d.bind(".", "q"); d.bind("a.b.c.d.e.f.g.h.i", "q")
    assert d.stats()["wstride"] == 4 and d.stats()["words"] == 2
    f = GpuMatcher(CHQ_FANOUT, device=-1)
    f.bind("x", "q1"); f.bind("y", "q2")
    assert f.stats()["groups"] == 1


def test_generator_is_deterministic_and_compiles_at_scale():
    w1, w2 = TopicWorkload(2), TopicWorkload(2)
    b1, o1 = w1.keys(5000, stream=3)
    b2, o2 = w2.keys(5000, stream=3)
    assert np.array_equal(b1, b2) and np.array_equal(o1, o2)
    b3, _ = w1.keys(5000, stream=4)
    assert not np.array_equal(b1[:1000], b3[:1000])
    pb, poff, q, d = w1.patterns(50_000, 12_500, hash_tail=True)
    assert q.max() == 12_499 and set(np.unique(d)) == {0, 1}
    qb, qoff = queue_names_packed(q)
    assert bytes(qb[:8]) == b"q0000000" and bytes(qb[-8:]) == b"q%07d" % q[-1]
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_SPEC, device=-1)
    m.bind_many(pb, poff, qb, qoff, d)
    m.commit()
    st = m.stats()
    assert st["queues"] == 12_500 and 40_000 < st["groups"] <= st["bindings"] <= 50_000
    assert st["shapes"] < 120 and st["wstride"] == 8


def test_three_byte_ids_host_side():  # chq_set_ids_format / chq_unpack_u24: the host half needs no GPU
    import numpy as np
    import pytest

    from chanamq_b200 import ChqError
    from chanamq_b200._lib import CHQ_IDS_U24, CHQ_IDS_U32
    m = GpuMatcher(CHQ_TOPIC, CHQ_MODE_COMPAT, device=-1)
    m.set_ids_format(CHQ_IDS_U24)
    m.set_ids_format(CHQ_IDS_U32)
    with pytest.raises(ChqError):
        m.set_ids_format(2)
    ids = np.array([0, 1, 255, 256, 65535, 65536, (1 << 24) - 1, 0x123456, 7], dtype=np.uint32)
    packed = np.zeros(3 * len(ids), dtype=np.uint8)
    for j, v in enumerate(ids.tolist()):  # the layout include/chanamq_gpu.h states: little-endian, back to back
        packed[3 * j], packed[3 * j + 1], packed[3 * j + 2] = v & 255, (v >> 8) & 255, v >> 16
    assert np.array_equal(m.unpack_u24(packed), ids)
    assert len(m.unpack_u24(np.zeros(0, np.uint8))) == 0
    m.close()


def test_header_is_plain_c():
    """The boundary is a C ABI: include/*.h must compile as C99 on its own (what a cgo / JNI / JNA generator is fed)."""
    import glob
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for h in sorted(glob.glob(os.path.join(root, "include", "*.h"))):
        out = subprocess.run(["gcc", "-x", "c", "-std=c99", "-fsyntax-only", "-Wall", "-Wextra", "-pedantic", "-Werror", h],
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stderr


def test_c99_consumer_links_and_runs(tmp_path):
    """tests/c_consumer.c: the host half of the boundary driven from plain C through the shared library, no Python in between"""
    import os
    import subprocess

    from chanamq_b200 import build

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.dirname(build.build_product())
    exe = str(tmp_path / "c_consumer")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(root, "include"),
                    os.path.join(root, "tests", "c_consumer.c"), "-o", exe, "-L", lib_dir, "-lchanagpu", "-Wl,-rpath," + lib_dir],
                   check=True, capture_output=True, text=True)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stdout + out.stderr
