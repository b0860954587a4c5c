"""Spec 4 ("classed sparse"; the default): an event with p < 2^-5 works on groups of 2^k samples,
k = min(13, z + 2) by its leading zero bits z; per group ONE Philox call gives the number of ones (cumulative
Binomial(2^k, thr/2^64) table) and their positions (k-bit fields); every other event samples as spec 2.  CPU tests pin the
oracle (class rule and table against big-integer and exact binomial arithmetic, streaming form against the definition, law
of the bits); GPU tests are bit-exact parity."""
import ctypes as C
import os
from fractions import Fraction
from math import comb

import numpy as np
import pytest

import spec_restatement as R
from helpers import G, O, eval_lowered, random_graph
from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan

CORES = len(os.sched_getaffinity(0))
K = 32                  # CANOPY_SPARSE_CDF_LEN
P_SPARSE = 2.0 ** -5    # 2^-CANOPY_SPARSE_MIN_ZEROS: events below it take the sparse path
# 64-bit thresholds across the classes: (thr64, k)
TABLE_CASES = [(1, 13), (5 << 32, 13), (429496 << 32, 13), ((1 << 52) - 1, 13), (1 << 52, 13), ((1 << 53) - 1, 13), (1 << 53, 12),
               (3 << 53, 11), (12345678 << 32, 10), (1 << 57, 8), ((1 << 58) - 1, 8), (1 << 58, 7), ((1 << 59) - 1, 7),
               (0x0123456789ABCDEF >> 4, 13), (0x0123456789ABCDEF << 1, 8)]
# probabilities covering dense, the boundary, and every class k = 7 .. 13
CLASS_PS = [1e-9, 3e-5, 1.1e-4, 2.0 ** -12, 4e-4, 2.0 ** -11, 9e-4, 0.0018, 2.0 ** -8, 0.0039, 0.0078, 0.011, 0.02, P_SPARSE - 2.0 ** -31,
            P_SPARSE, 0.1, 0.5, 2.0 ** -31]


def _lib_cdf(t, k):
    out = (C.c_uint64 * K)()
    _lib.check(_lib.lib().canopy_sparse_cdf(t, k, out))
    return [int(x) for x in out]


def _lib_class(t):
    k = C.c_uint32()
    _lib.check(_lib.lib().canopy_sparse_class(t, C.byref(k)))
    return int(k.value)


def _popcount(w):
    return sum(bin(int(x)).count("1") for x in np.asarray(w).ravel())


def _graph_of(ps):
    b = G.GraphBuilder()
    b.root(b.gate(G.OR, [b.basic_event(p) for p in ps]))
    return b.build()


@pytest.mark.parametrize("t,k", TABLE_CASES)
def test_class_and_count_table_three_ways(t, k):
    """The library's class and table, the oracle's and the Python big-integer restatement agree word for word, and the
    table is the Binomial(2^k, t/2^64) distribution function to better than 2^-50."""
    assert _lib_class(t) == O.sparse_class_k(t) == R.class_k(t) == k
    lib = _lib_cdf(t, k)
    assert lib == O.sparse_cdf(t, k) == list(R.count_table(t, k))
    assert all(a <= b for a, b in zip(lib, lib[1:]))
    p, acc, n = Fraction(t, 1 << 64), Fraction(0), 1 << k
    q_pow = (1 - p) ** n
    term = q_pow
    for j in range(K):
        acc += term
        assert abs(acc - Fraction(lib[j], 1 << 64)) < Fraction(1, 1 << 50), j
        term = term * (n - j) * p / ((j + 1) * (1 - p))
    assert 1 - acc < Fraction(1, 10 ** 17)       # mass beyond 32 ones, folded into "32" (largest for k = 13 at four expected ones)


def test_class_rule_edges():
    assert [_lib_class(x) for x in (0, 1 << 59, (1 << 59) - 1, 1 << 63, (1 << 64) - 1)] == [0, 0, 7, 0, 0]
    for z in range(5, 64):
        t = 1 << (63 - z)                         # exactly z leading zeros
        assert _lib_class(t) == O.sparse_class_k(t) == min(13, z + 2)
        assert (1 << _lib_class(t)) * Fraction(2 * t - 1, 1 << 64) < 4      # fewer than four expected ones per group


def test_count_table_rejects_dense_thresholds_and_bad_k():
    out = (C.c_uint64 * K)()
    f = _lib.lib().canopy_sparse_cdf
    assert f(1 << 59, 7, out) == 1 and f(0, 7, out) == 1 and f(1 << 40, 6, out) == 1 and f(1 << 40, 14, out) == 1


@pytest.mark.parametrize("bits", [32, 64])
def test_streaming_form_equals_the_definition(bits):
    """Sparse events of every class, the boundary, and dense events (which must be spec 2 bit for bit)."""
    g = _graph_of(CLASS_PS)
    for seed, first in ((0, 0), ((1 << 63) | 77, (1 << 40) + 126)):       # the second range straddles an 8192-sample boundary
        slow = O.sample_words(g, seed, range(len(CLASS_PS)), first, 6, bits, 4, fast=False)
        fast = O.sample_words(g, seed, range(len(CLASS_PS)), first, 6, bits, 4, fast=True)
        assert np.array_equal(slow, fast)
    seed, first = 5, 64 * 9
    fast = O.sample_words(g, seed, range(len(CLASS_PS)), first, 256, bits, 4, fast=True)
    spec2 = O.sample_words(g, seed, range(len(CLASS_PS)), first, 256, bits, 2, fast=True)
    for i, p in enumerate(CLASS_PS):
        assert np.array_equal(spec2[i], fast[i]) == (p >= P_SPARSE or _popcount(fast[i]) + _popcount(spec2[i]) == 0), p


def test_streaming_form_equals_the_definition_on_random_inputs():
    """Property test over random thresholds (log-uniform across the sparse range and just above it), seeds, events, windows."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=150, deadline=None, derandomize=True)
    @given(st.floats(min_value=-40.0, max_value=-4.5), st.integers(0, 2 ** 64 - 1), st.integers(0, 2 ** 32 - 1),
           st.integers(0, 2 ** 57 - 1), st.sampled_from([32, 64]))
    def check(log2p, seed, e_pad, win, bits):
        b = G.GraphBuilder()
        ev = [b.basic_event(2.0 ** log2p) for _ in range(1 + e_pad % 5)]
        b.root(b.gate(G.NULL, [ev[-1]]))
        g = b.build()
        slow = O.sample_words(g, seed, [len(ev) - 1], 2 * win, 2, bits, 4, fast=False)
        fast = O.sample_words(g, seed, [len(ev) - 1], 2 * win, 2, bits, 4, fast=True)
        assert np.array_equal(slow, fast)

    check()


def _forced_tables():
    full = np.zeros(K, dtype=np.uint64)                                     # every group draws 32 ones
    stair = np.array([0] * 20 + [1 << 63] * 12, dtype=np.uint64)            # 20 or 32 ones per group
    return {"all32": full, "20or32": stair}


@pytest.mark.parametrize("k", [7, 8, 10, 11, 13])
@pytest.mark.parametrize("name", ["all32", "20or32"])
def test_forced_counts_reach_the_extension_calls(name, k):
    """With the count table overridden (test hook) every group needs 20 to 40+ position fields, i.e. calls 0x401, 0x402
    and beyond (4, 3 or 2 fields per word for k = 7-8, 9-10, 11-13): the two oracle forms agree there too."""
    tab = _forced_tables()[name]
    n_words = (1 << k) // 64 * 6
    a = O.sparse_words_with_table(2024, 7, k, tab, 128 * 5, n_words, stream=True)
    b = O.sparse_words_with_table(2024, 7, k, tab, 128 * 5, n_words, stream=False)
    assert np.array_equal(a, b)
    first_group_word = -(-128 * 5 * 64 // (1 << k)) * (1 << k) // 64 - 128 * 5      # first whole group inside the range
    whole = a[first_group_word: first_group_word + (1 << k) // 64 * 4]
    per_group = np.unpackbits(whole.view(np.uint8)).reshape(-1, 1 << k).sum(axis=1)
    assert set(per_group.tolist()) <= ({32} if name == "all32" else {20, 32}) and per_group.max() == 32


def test_law_of_the_sparse_bits():
    """Frequencies within 4 sigma of thr/2^32; ones per group distributed as Binomial(2^k, p); positions uniform; neighbours
    uncorrelated -- for one event of each of four classes."""
    ps = [1.5e-4, 1.7e-3, 2.0 ** -8 * 0.999, 0.03]         # k = 13, 10, 9, 7
    b = G.GraphBuilder()
    for p in ps:
        b.root(b.gate(G.NULL, [b.basic_event(p)]))
    g = b.build()
    n = 1 << 24
    t = O.mc_packed(g, 99, 12345, n, nthreads=CORES, spec=4).astype(np.float64)
    q = np.array([O.quantised_p(p) for p in ps])
    assert np.all(np.abs(t / n - q) <= 4.0 * np.sqrt(q * (1 - q) / n))
    for idx, kmax in ((0, 4), (1, 6), (2, 6), (3, 10)):
        k = O.sparse_class_k(int(q[idx] * 2 ** 64))
        n_grp = (1 << 24) >> k
        w = O.sample_words(g, 7, [idx], 0, n_grp << (k - 6), 32, 4, fast=True)[0]
        bits = np.unpackbits(w.view(np.uint8), bitorder="little").reshape(n_grp, 1 << k)
        counts = bits.sum(axis=1)
        for c in range(kmax):
            pc = comb(1 << k, c) * q[idx] ** c * (1 - q[idx]) ** ((1 << k) - c)
            assert abs((counts == c).sum() - n_grp * pc) <= 4.5 * np.sqrt(n_grp * pc * (1 - pc)) + 1, (idx, c)
        per_pos = bits.sum(axis=0).astype(np.float64)
        assert np.all(np.abs(per_pos - n_grp * q[idx]) <= 5.0 * np.sqrt(n_grp * q[idx]) + 1)
        both = int((bits[:, :-1] & bits[:, 1:]).sum())
        ex = n_grp * ((1 << k) - 1) * q[idx] ** 2
        assert abs(both - ex) <= 4.5 * np.sqrt(ex) + 2


@pytest.mark.parametrize("seed", range(4))
def test_mc_within_4_sigma_of_exact_enumeration(seed):
    g = random_graph(seed + 900, n_basic=10, n_gates=14, n_roots=2, extreme_p=False, consts=False, p_lo=0.0005, p_hi=0.031)
    exact = O.exact(g, True)
    n = 1 << 20
    t = O.mc_packed(g, seed, 0, n, True, nthreads=CORES, spec=4).astype(np.float64)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / n)
    assert np.all(np.abs(t / n - exact) <= 4.0 * sigma + 3.0 / n)   # + 3 counts: gates with n p << 1 are Poisson, not normal


def test_events_are_independent_of_each_other():
    """AND of two or three sparse events (of different classes) fires with the product of their probabilities."""
    b = G.GraphBuilder()
    ps = [0.02, 0.03, 0.0078, 0.001, 0.031, 0.015]
    e = [b.basic_event(p) for p in ps]
    pairs = [(0, 1), (0, 2), (1, 4), (2, 3), (4, 5), (0, 5)]
    for i, j in pairs:
        b.root(b.gate(G.AND, [e[i], e[j]]))
    b.root(b.gate(G.AND, [e[0], e[1], e[4]]))
    n = 1 << 24
    t = O.mc_packed(b.build(), 4242, 0, n, nthreads=CORES, spec=4).astype(np.float64)
    q = [O.quantised_p(p) for p in ps]
    ex = np.array([q[i] * q[j] for i, j in pairs] + [q[0] * q[1] * q[4]])
    assert np.all(np.abs(t / n - ex) <= 4.0 * np.sqrt(ex * (1 - ex) / n))


@pytest.mark.parametrize("bits", [32, 64])
def test_scalar_packed_and_lowered_agree_under_spec4(bits):
    g = random_graph(31, n_basic=14, n_gates=30, n_roots=3, p_lo=1e-4, p_hi=0.02)
    begin, n = 8192 * 2 - 1500, 3000
    want = O.mc_scalar(g, 5, begin, n, True, bits, spec=4)
    assert np.array_equal(O.mc_packed(g, 5, begin, n, True, bits, nthreads=3, spec=4), want)
    assert np.array_equal(O.mc_packed(g, 5, begin, 1000, True, bits, spec=4) + O.mc_packed(g, 5, begin + 1000, 2000, True, bits, spec=4), want)
    with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=4) as plan:     # host lowering only: no GPU needed
        assert plan.spec == 4
        v = plan.view()
        assert np.array_equal(eval_lowered(v, g, 5, begin, n, bits, spec=4), want)
        assert [int(k) for k in v["event_k"]] == [O.sparse_class_k(int(t)) for t in v["event_thr64"]]


def test_spec3_is_retired():
    with pytest.raises(_lib.CanopyError) as ei:
        Plan(random_graph(1), sampler=3)
    assert ei.value.status == _lib.E_UNSUPPORTED


# ---------------------------------------------------------------------------------------------------------------
# GPU parity
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
def test_gpu_sampler_words_every_live_event(bits):
    g = synth.config_graph("c2", 0.05)
    with Plan(g, threshold_bits=bits, sampler=4) as plan:
        v = plan.view()
        got = plan.sample_words(seed=(1 << 63) + 12345, first_word=64 * 1000, n_words=256)
    assert np.array_equal(got, O.sample_words(g, (1 << 63) + 12345, v["event_orig"], 64 * 1000, 256, bits, 4, fast=True))


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
def test_gpu_every_class_and_the_boundary_against_the_definition(bits):
    g = _graph_of(CLASS_PS)
    with Plan(g, threshold_bits=bits, sampler=4) as plan:
        v = plan.view()
        got = plan.sample_words(seed=77, first_word=64 * 3, n_words=128)
    assert np.array_equal(got, O.sample_words(g, 77, v["event_orig"], 64 * 3, 128, bits, 4, fast=False))
    assert sorted(set(int(k) for k in v["event_k"])) == [0] + list(range(7, 14))


@pytest.mark.gpu
def test_gpu_units_with_fewer_than_four_events_and_many_classes():
    """1, 2, 3, 5 and 9 events per class: partial units at the end of every class, units of one class only."""
    ps = []
    for k, cnt in zip(range(7, 14), (1, 2, 3, 5, 9, 1, 6)):
        ps += [2.0 ** -(k - 2) * (0.51 + 0.03 * i) for i in range(cnt)]
    ps += [0.2, 0.04]
    g = _graph_of(ps)
    with Plan(g, sampler=4) as plan:
        v = plan.view()
        assert [int(x) for x in np.bincount(v["event_k"], minlength=14)[7:]] == [1, 2, 3, 5, 9, 1, 6]
        got = plan.sample_words(seed=3, first_word=128 * 40, n_words=128 * 5)
    assert np.array_equal(got, O.sample_words(g, 3, v["event_orig"], 128 * 40, 128 * 5, 32, 4, fast=True))


@pytest.mark.gpu
def test_gpu_long_sparse_streams_match_oracle():
    """2^27 samples of events of every class: every count the tables give and many duplicate fields."""
    ps = [0.03, 0.0155, 0.0078, 0.0039, 0.0019, 0.00097, 0.00048, 0.0002, 0.031, 0.0305]
    g = _graph_of(ps)
    n_words = 1 << 21
    with Plan(g, sampler=4) as plan:
        got = plan.sample_words(seed=2024, first_word=0, n_words=n_words)
    want = O.sample_words(g, 2024, range(len(ps)), 0, n_words, 32, 4, fast=True)
    assert np.array_equal(got, want)
    per_group = np.unpackbits(want[0].view(np.uint8)).reshape(-1, 128).sum(axis=1)
    assert per_group.max() >= 14 and (per_group > 8).mean() > 0.01        # the second position call is routine at p = 0.03


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["all32", "20or32"])
def test_gpu_forced_counts_reach_the_extension_calls(name):
    """Kernel against oracle where every group consumes 20 to 40+ position fields (count table overridden by the test hook),
    for one event of each field width (k = 7, 10, 13)."""
    tab = _forced_tables()[name]
    ps = (0.5, 0.01, 0.02, 0.2, 0.003, 0.0001)
    b = G.GraphBuilder()
    ev = [b.basic_event(p) for p in ps]
    b.root(b.gate(G.OR, ev))
    g = b.build()
    n_words = 128 * 64
    with Plan(g, sampler=4) as plan:
        orig = [int(x) for x in plan.view()["event_orig"]]
        ks = {int(e): int(k) for e, k in zip(plan.view()["event_orig"], plan.view()["event_k"])}
        assert (ks[2], ks[4], ks[5]) == (7, 10, 13)
        before = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
        for e in (2, 4, 5):
            plan.debug_set_sparse_table(orig.index(e), tab)
        got = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
        with pytest.raises(Exception):
            plan.debug_set_sparse_table(orig.index(0), tab)          # p = 0.5 is not a sparse event
        plan.update_p(list(ps))                                      # restores the real tables
        after = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
    want = O.sample_words(g, 2024, orig, 128 * 5, n_words, 32, 4, fast=True)
    assert np.array_equal(before, want) and np.array_equal(after, want)
    for e in (2, 4, 5):
        want[orig.index(e)] = O.sparse_words_with_table(2024, e, ks[e], tab, 128 * 5, n_words, stream=True)
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(6))
def test_gpu_random_graphs_all_connectives(seed):
    g = random_graph(300 + seed, n_basic=16, n_gates=48, max_fanin=9, n_roots=4, p_lo=1e-4, p_hi=0.05)
    bits = 64 if seed % 2 else 32
    with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=4) as plan:
        got, _ = plan.run(20_000, seed=seed, sample_begin=8192 * seed + 7000)
    assert np.array_equal(got, O.mc_scalar(g, seed, 8192 * seed + 7000, 20_000, True, bits, spec=4))


@pytest.mark.gpu
@pytest.mark.parametrize("name,scale,n", [("c2", 0.1, 1 << 15), ("c3", 0.05, 1 << 15), ("c4", 0.02, 1 << 14), ("c2", 1.0, 1 << 13)])
def test_gpu_configs(name, scale, n):
    g = synth.config_graph(name, scale)
    begin = (1 << 40) + 4096 * 9 + 13
    with Plan(g, tally_all_gates=True, sampler=4) as plan:
        got, rep = plan.run(n, seed=77, sample_begin=begin)
        halves = plan.run(n // 3, seed=77, sample_begin=begin)[0] + plan.run(n - n // 3, seed=77, sample_begin=begin + n // 3)[0]
    assert np.array_equal(got, O.mc_packed(g, 77, begin, n, True, nthreads=CORES, spec=4))
    assert np.array_equal(got, halves)


@pytest.mark.gpu
def test_gpu_64_bit_thresholds_resolve_what_32_bits_cannot():
    """p = 1e-8 and below: a 32-bit threshold is off by percents (or zero), a 64-bit one is exact to 2^-64; both run the
    sparse path (k = 13) and match the oracle; the 64-bit frequencies match the probabilities."""
    ps = [1e-8, 3e-10, 2.0 ** -33, 1e-6]
    b = G.GraphBuilder()
    for p in ps:
        b.root(b.gate(G.NULL, [b.basic_event(p)]))
    g = b.build()
    n = 1 << 30
    for bits in (32, 64):
        with Plan(g, threshold_bits=bits, sampler=4) as plan:
            got, _ = plan.run(n, seed=8)
            short, _ = plan.run(1 << 22, seed=8, sample_begin=1 << 29)
            live = {int(e): int(t) for e, t in zip(plan.view()["event_orig"], plan.view()["event_thr64"])}
        assert np.array_equal(short, O.mc_packed(g, 8, 1 << 29, 1 << 22, False, bits, nthreads=CORES, spec=4))
        q = np.array([O.quantised_p(p, bits) for p in ps])
        assert np.all(np.abs(got - n * q) <= 4.5 * np.sqrt(n * q) + 1)
        if bits == 64:
            # 2^-33 is sampled with its exact threshold (expected count over 2^30 samples: 1/8, so no count is asserted)
            assert np.all(np.abs(q - np.array(ps)) <= 1e-18) and live[2] == 1 << 31
        else:
            assert int(got[2]) == 0 and 2 not in live      # ... and can never fire with 32-bit thresholds: folded to a constant


@pytest.mark.gpu
def test_gpu_update_p_moves_events_across_classes():
    b = G.GraphBuilder()
    e = [b.basic_event(p) for p in (0.001, 0.05, 0.004, 0.3)]
    b.root(b.gate(G.OR, e[:2]))
    b.root(b.gate(G.ATLEAST, e, 2))
    g = b.build()
    n = 1 << 16
    with Plan(g, tally_all_gates=True, sampler=4) as plan:
        first, _ = plan.run(n, seed=11)
        plan.update_p([0.02, 0.0005, 0.004, 0.006])
        second, _ = plan.run(n, seed=11)
    g2 = G.from_lists([0.02, 0.0005, 0.004, 0.006], [(G.OR, [0, 1], 0, 0), (G.ATLEAST, [0, 1, 2, 3], 2, 0)], [4, 5])
    assert np.array_equal(first, O.mc_packed(g, 11, 0, n, True, nthreads=CORES, spec=4))
    assert np.array_equal(second, O.mc_packed(g2, 11, 0, n, True, nthreads=CORES, spec=4))


@pytest.mark.gpu
def test_gpu_forced_counts_in_partly_filled_units():
    """Units whose groups do not fill the warp (three k = 13 events: 3 of 32 lanes busy; two k = 11 events: 8 of 32) with an
    all-zero count table: the idle lanes compare U = 0 with cdf_0 = 0 and would draw 32 ones into row 0 if the kernel did
    not switch them off explicitly.  Every row must equal the oracle's for that event alone."""
    tab = _forced_tables()["all32"]
    ps = (0.5, 0.0001, 0.00011, 0.00012, 0.001, 0.0015, 0.02)
    b = G.GraphBuilder()
    ev = [b.basic_event(p) for p in ps]
    b.root(b.gate(G.OR, ev))
    g = b.build()
    n_words = 128 * 64
    with Plan(g, sampler=4) as plan:
        orig = [int(x) for x in plan.view()["event_orig"]]
        ks = {int(e): int(k) for e, k in zip(plan.view()["event_orig"], plan.view()["event_k"])}
        assert [ks[e] for e in range(1, 7)] == [13, 13, 13, 11, 11, 7]
        for e in range(1, 7):
            plan.debug_set_sparse_table(orig.index(e), tab)
        got = plan.sample_words(seed=77, first_word=128 * 3, n_words=n_words)
    want = O.sample_words(g, 77, orig, 128 * 3, n_words, 32, 4, fast=True)
    for e in range(1, 7):
        want[orig.index(e)] = O.sparse_words_with_table(77, e, ks[e], tab, 128 * 3, n_words, stream=True)
    assert np.array_equal(got, want)
