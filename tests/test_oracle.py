"""Oracle pinned against the only external vectors that exist for this path (Philox KATs), against exact
enumeration, and against the committed golden fixtures.  CPU only."""
import json
import math
import os

import numpy as np
import pytest

from helpers import G, O, random_graph

HERE = os.path.dirname(os.path.abspath(__file__))


def load_golden():
    with open(os.path.join(HERE, "golden", "golden_tallies.json")) as f:
        return json.load(f)["cases"]


def graph_of(case):
    g = case["graph"]
    return G.Graph(p=g["p"], op=g["op"], kmin=g["kmin"], kmax=g["kmax"], child_off=g["child_off"], child=g["child"],
                   roots=g["roots"])


def test_philox_known_answer_vectors():
    with open(os.path.join(HERE, "golden", "philox_kat.json")) as f:
        for v in json.load(f)["vectors"]:
            assert O.philox4x32_10(v["ctr"], v["key"]) == v["out"]


def test_threshold_quantisation():
    assert O.threshold(0.0) == (0, 0)
    assert O.threshold(1.0) == (1, 0)
    assert O.threshold(0.5) == (2, 1 << 31)
    assert O.threshold(0.5, 64) == (2, 1 << 63)
    assert O.threshold(2.0 ** -32) == (2, 1)
    assert O.threshold(2.0 ** -33) == (2, 0)      # below resolution: never fires (documented in DESIGN.md)
    assert O.threshold(2.0 ** -33, 64)[1] == 1 << 31
    assert O.threshold(float("nan"))[0] == -1
    assert O.threshold(1.5)[0] == -1 and O.threshold(-0.1)[0] == -1
    assert O.quantised_p(0.1) == math.floor(0.1 * 2 ** 32) / 2 ** 32


@pytest.mark.parametrize("case", load_golden(), ids=lambda c: c["name"])
def test_golden_scalar_and_packed(case):
    g = graph_of(case)
    want = np.array(case["tallies"], dtype=np.uint64)
    spec = case["spec"]
    got = O.mc_scalar(g, case["seed"], case["begin"], case["n"], case["tally_all"], case["bits"], spec)
    assert np.array_equal(got, want)
    got_p = O.mc_packed(g, case["seed"], case["begin"], case["n"], case["tally_all"], case["bits"], nthreads=3, spec=spec)
    assert np.array_equal(got_p, want)
    for fast in (False, True):
        words = O.sample_words(g, case["seed"], case["word_events"], case["first_word"], 2, case["bits"], spec, fast)
        assert np.array_equal(words, np.array(case["words"], dtype=np.uint64).reshape(words.shape))


@pytest.mark.parametrize("spec", [1, 2, 4])
@pytest.mark.parametrize("bits", [32, 64])
def test_lazy_group_sampler_equals_the_written_definition(spec, bits):
    """mc_scalar / mc_packed take basic-event bits from the lazily evaluated 128-sample group routine; the
    definition (oracle_sample_bit) builds each variate in full. Same words, including p whose top threshold
    byte is 0x00 / 0xFF / mixed, and tails that are all-zero."""
    b = G.GraphBuilder()
    ps = [1e-9, 3e-5, 2.0 ** -8, 2.0 ** -8 + 2.0 ** -30, 0.0039, 0.1, 0.5, 0.75, 1 - 2.0 ** -9, 0.999999, 2.0 ** -31]
    e = [b.basic_event(p) for p in ps]
    b.root(b.gate(G.OR, e))
    g = b.build()
    for seed, first in ((0, 0), ((1 << 63) | 77, (1 << 40) + 6)):
        slow = O.sample_words(g, seed, range(len(ps)), first, 8, bits, spec, fast=False)
        fast = O.sample_words(g, seed, range(len(ps)), first, 8, bits, spec, fast=True)
        assert np.array_equal(slow, fast)
    # and the single-bit entry point agrees with the packed words
    w = O.sample_words(g, 5, [5], 3, 1, bits, spec)[0, 0]
    assert [O.sample_bit(5, 5, 3 * 64 + i, ps[5], bits, spec) for i in range(64)] == [(int(w) >> i) & 1 for i in range(64)]


@pytest.mark.parametrize("spec", [1, 2, 4])
def test_sampled_frequency_matches_quantised_probability(spec):
    """Per-event frequency over 2^20 samples within 4 sigma of thr/2^32, for thresholds whose top byte is zero,
    mixed and all-ones (exercises head-only, head+tail and tail-dominated decisions of the sliced spec)."""
    b = G.GraphBuilder()
    ps = [1e-4, 2.0 ** -8 * 0.999, 0.0123, 0.3, 0.5, 0.9, 0.998]
    e = [b.basic_event(p) for p in ps]
    for x in e:
        b.root(b.gate(G.NULL, [x]))
    g = b.build()
    n = 1 << 20
    t = O.mc_packed(g, 99, 12345, n, nthreads=4, spec=spec).astype(np.float64)
    q = np.array([O.quantised_p(p) for p in ps])
    assert np.all(np.abs(t / n - q) <= 4.0 * np.sqrt(q * (1 - q) / n))


@pytest.mark.parametrize("spec", [1, 2, 4])
@pytest.mark.parametrize("seed", range(5))
def test_mc_within_3_sigma_of_exact_enumeration(seed, spec):
    g = random_graph(seed + 50, n_basic=10, n_gates=16, n_roots=2, extreme_p=False)
    exact = O.exact(g, True)
    n = 1 << 16
    t = O.mc_packed(g, seed, 0, n, True, nthreads=4, spec=spec).astype(np.float64)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / n)
    assert np.all(np.abs(t / n - exact) <= 4.0 * sigma + 1e-9)


def test_range_split_invariance():
    g = random_graph(7)
    whole = O.mc_packed(g, 3, 10, 9000, True, nthreads=2)
    parts = O.mc_packed(g, 3, 10, 4000, True) + O.mc_packed(g, 3, 4010, 5000, True, nthreads=3)
    assert np.array_equal(whole, parts)


def test_connective_truth_tables():
    """Each connective against its definition through exact enumeration at p = 1/2."""
    b = G.GraphBuilder()
    e = [b.basic_event(0.5) for _ in range(4)]
    cases = {
        "and": (G.AND, e[:3], 0, 0, lambda x: all(x[:3])),
        "or": (G.OR, e[:3], 0, 0, lambda x: any(x[:3])),
        "atleast2of4": (G.ATLEAST, e, 2, 0, lambda x: sum(x) >= 2),
        "xor": (G.XOR, e[:2], 0, 0, lambda x: x[0] != x[1]),
        "not": (G.NOT, e[:1], 0, 0, lambda x: not x[0]),
        "nand": (G.NAND, e[:3], 0, 0, lambda x: not all(x[:3])),
        "nor": (G.NOR, e[:3], 0, 0, lambda x: not any(x[:3])),
        "null": (G.NULL, e[:1], 0, 0, lambda x: x[0]),
        "iff": (G.IFF, e[:2], 0, 0, lambda x: x[0] == x[1]),
        "imply": (G.IMPLY, e[:2], 0, 0, lambda x: (not x[0]) or x[1]),
        "card1to2of4": (G.CARDINALITY, e, 1, 2, lambda x: 1 <= sum(x) <= 2),
    }
    for name, (op, args, lo, hi, fn) in cases.items():
        b.root(b.gate(op, args, lo, hi, name=name))
    probs = O.exact(b.build())
    for i, (name, (_, _, _, _, fn)) in enumerate(cases.items()):
        want = sum(fn([(s >> k) & 1 for k in range(4)]) for s in range(16)) / 16
        assert abs(probs[i] - want) < 1e-12, name


def test_cycle_is_reported():
    b = G.GraphBuilder()
    e = b.basic_event(0.1)
    g0 = b.gate(G.OR, [e, e])  # placeholder, redefined below
    g1 = b.gate(G.AND, [g0, e])
    b.set_gate(g0, G.OR, [g1, e])
    b.root(g1)
    with pytest.raises(ValueError, match="cycle"):
        O.mc_scalar(b.build(), 0, 0, 10)


def test_vectorised_philox_of_the_cpu_port_equals_the_scalar_one():
    import random
    rng = random.Random(3)
    for _ in range(50):
        a = [rng.getrandbits(32) for _ in range(3)]
        c3 = [rng.getrandbits(32) for _ in range(8)]
        key = [rng.getrandbits(32), rng.getrandbits(32)]
        got = O.philox_batch8(a[0], a[1], a[2], c3, key)
        for i in range(8):
            assert got[i] == O.philox4x32_10([a[0], a[1], a[2], c3[i]], key)
