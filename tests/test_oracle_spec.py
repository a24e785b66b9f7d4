"""SPEC-mode oracle (AMQP 0-9-1 '*'/'#'; parity unpinned — the reference has no '#', QueueMatcher.scala:70).
Three opinions must agree: the DP predicate, the SpecTrie index, and — on '#'-free tables — the
reference-exact COMPAT trie."""
import random

from oracle import pyoracle as po


def test_amqp_spec_example():  # SURVEY Appendix C KAT
    assert po.pair_match(po.SPEC, "*.stock.#", "usd.stock") == 1
    assert po.pair_match(po.SPEC, "*.stock.#", "eur.stock.db") == 1
    assert po.pair_match(po.SPEC, "*.stock.#", "stock.nasdaq") == 0
    assert po.pair_match(po.SPEC, "#", "a.b.c") == 1
    assert po.pair_match(po.SPEC, "#.c", "c") == 1
    assert po.pair_match(po.SPEC, "a.#.#.b", "a.b") == 1
    assert po.pair_match(po.SPEC, "a.#.b.#", "a.x.y.b.z") == 1
    assert po.pair_match(po.SPEC, "a.#.b", "a.b.c") == 0
    assert po.pair_match(po.SPEC, "*.#", "") == 1  # one (empty) word
    assert po.pair_match(po.COMPAT, "a.#", "a.b") == 0 and po.pair_match(po.COMPAT, "a.#", "a.#") == 1


def _rand_topic(rng, alphabet, maxw, wild):
    n = rng.randint(1, maxw)
    return ".".join(rng.choice(wild) if rng.random() < 0.35 else rng.choice(alphabet) for _ in range(n))


def test_spec_trie_agrees_with_dp_predicate():
    rng = random.Random(7)
    alpha = ["a", "b", "c", "", "dd"]
    for trial in range(30):
        pats = [_rand_topic(rng, alpha, 5, ["*", "#", "#"]) for _ in range(40)]
        keys = [_rand_topic(rng, alpha + ["*", "#"], 6, ["x"]) for _ in range(60)]
        o = po.Oracle(po.TOPIC, po.SPEC)
        live = []
        for i, p in enumerate(pats):
            if o.bind(p, "q%03d" % (i % 25)) == 0:
                live.append((p, "q%03d" % (i % 25)))
        got = o.route(keys)
        for k, g in zip(keys, got):
            if g is None:
                assert po.split(k) == []
                continue
            want = sorted({q.encode() for p, q in live if po.pair_match(po.SPEC, p, k) == 1})
            assert g == want, (k, pats)


def test_compat_trie_agrees_with_predicate_and_with_spec_on_hash_free_tables():
    rng = random.Random(11)
    alpha = ["a", "b", "c", "", "dd", "#x"]
    for trial in range(30):
        pats = [_rand_topic(rng, alpha, 5, ["*"]) for _ in range(50)]
        keys = [_rand_topic(rng, alpha + ["*"], 5, ["y"]) for _ in range(80)]
        oc, os_ = po.Oracle(po.TOPIC, po.COMPAT), po.Oracle(po.TOPIC, po.SPEC)
        live = []
        for i, p in enumerate(pats):
            q = "q%03d" % (i % 20)
            rc = oc.bind(p, q)
            assert rc == os_.bind(p, q)
            if rc == 0:
                live.append((p, q))
        rc_, rs_ = oc.route(keys), os_.route(keys)
        assert rc_ == rs_
        for k, g in zip(keys, rc_):
            if g is not None:
                assert g == sorted({q.encode() for p, q in live if po.pair_match(po.COMPAT, p, k) == 1})
