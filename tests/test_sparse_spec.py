"""Spec 3 ("sparse"; the default): events with fewer than four expected ones per 128-sample group draw the
group's number of ones from a cumulative Binomial(128, thr/2^32) table and place them with 7-bit fields, all from ONE
Philox call; every other event samples as spec 2.  CPU tests pin the oracle (table against big-integer and exact
binomial arithmetic, streaming form against the definition, law of the bits); GPU tests are bit-exact parity."""
import ctypes as C
import os
from fractions import Fraction
from math import comb

import numpy as np
import pytest

from helpers import G, O, eval_lowered, random_graph
from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan

CORES = len(os.sched_getaffinity(0))
SPARSE_THRS = [1, 5, 429496, 1 << 20, 12345678, (1 << 25) - 1, 1 << 26, (1 << 27) - 1]
K = 32           # CANOPY_SPARSE_CDF_LEN
P_SPARSE = 2.0 ** -5   # 2^-CANOPY_SPARSE_MIN_ZEROS: events below it take the sparse path
# (event, group, ones) with 9 or more ones at seed 2024 among four events of p = 0.0078: the ninth position field is
# the first one of call 0x401 (found by scanning 3.1e6 groups with the oracle's streaming form)
EXT_GROUPS = [(1, 671172, 9), (3, 859388, 9), (2, 1926985, 9), (3, 3089365, 10)]


def _bigint_cdf(t):
    q = (1 << 32) - t
    x = q << 32
    for _ in range(7):
        x = (x * x) >> 64
    a, acc, out = x, 0, []
    for j in range(K):
        acc += a
        out.append(min(acc, (1 << 64) - 1))
        a = (a * (128 - j) * t) // ((j + 1) * q)
    return out


def _ext_graph():
    b = G.GraphBuilder()
    b.root(b.gate(G.OR, [b.basic_event(0.0078) for _ in range(4)]))
    return b.build()


def _popcount(w):
    return sum(bin(int(x)).count("1") for x in np.asarray(w).ravel())


@pytest.mark.parametrize("t", SPARSE_THRS)
def test_count_table_three_ways(t):
    """The library's table, the oracle's and a Python big-integer restatement are the same 32 words, and they are the
    Binomial(128, t/2^32) distribution function to better than 2^-55."""
    out = (C.c_uint64 * K)()
    _lib.check(_lib.lib().canopy_sparse_cdf(t, out))
    lib = [int(x) for x in out]
    assert lib == O.sparse_cdf(t) == _bigint_cdf(t)
    assert all(a <= b for a, b in zip(lib, lib[1:]))
    p, acc = Fraction(t, 1 << 32), Fraction(0)
    for j in range(K):
        acc += comb(128, j) * p ** j * (1 - p) ** (128 - j)
        assert abs(acc - Fraction(lib[j], 1 << 64)) < Fraction(1, 1 << 55)
    assert 1 - acc < Fraction(1, 10 ** 18)       # mass beyond 32 ones, folded into "32"


def test_count_table_rejects_dense_thresholds():
    out = (C.c_uint64 * K)()
    assert _lib.lib().canopy_sparse_cdf(1 << 27, out) == 1 and _lib.lib().canopy_sparse_cdf(0, out) == 1


def test_streaming_form_equals_the_definition():
    """Sparse events (p < 2^-5), the boundary, and dense events (which must be spec 2 bit for bit)."""
    b = G.GraphBuilder()
    ps = [1e-9, 3e-5, 2.0 ** -8, P_SPARSE - 2.0 ** -31, P_SPARSE, 0.0039, 0.1, 0.5, 2.0 ** -31, 1e-4, 0.0078, 0.02, 0.031]
    b.root(b.gate(G.OR, [b.basic_event(p) for p in ps]))
    g = b.build()
    for seed, first in ((0, 0), ((1 << 63) | 77, (1 << 40) + 6)):
        slow = O.sample_words(g, seed, range(len(ps)), first, 64, 32, 3, fast=False)
        fast = O.sample_words(g, seed, range(len(ps)), first, 64, 32, 3, fast=True)
        assert np.array_equal(slow, fast)
        spec2 = O.sample_words(g, seed, range(len(ps)), first, 64, 32, 2, fast=True)
        for i, p in enumerate(ps):
            assert np.array_equal(spec2[i], fast[i]) == (p >= P_SPARSE or _popcount(fast[i]) + _popcount(spec2[i]) == 0)
        # 64-bit thresholds: spec 3 is spec 2
        assert np.array_equal(O.sample_words(g, seed, range(len(ps)), first, 8, 64, 3, fast=True),
                              O.sample_words(g, seed, range(len(ps)), first, 8, 64, 2, fast=True))


@pytest.mark.parametrize("ev,grp,ones", EXT_GROUPS)
def test_groups_that_need_a_second_call(ev, grp, ones):
    g = _ext_graph()
    slow = O.sample_words(g, 2024, [ev], 2 * grp, 2, 32, 3, fast=False)
    fast = O.sample_words(g, 2024, [ev], 2 * grp, 2, 32, 3, fast=True)
    assert np.array_equal(slow, fast) and _popcount(fast) == ones


def test_streaming_form_equals_the_definition_on_random_inputs():
    """Property test over random thresholds (log-uniform across the sparse range and just above it), seeds, events, groups."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=150, deadline=None, derandomize=True)
    @given(st.floats(min_value=-31.9, max_value=-4.5), st.integers(0, 2 ** 64 - 1), st.integers(0, 2 ** 32 - 1),
           st.integers(0, 2 ** 57 - 1))
    def check(log2p, seed, e_pad, grp):
        b = G.GraphBuilder()
        ev = [b.basic_event(2.0 ** log2p) for _ in range(1 + e_pad % 5)]
        b.root(b.gate(G.NULL, [ev[-1]]))
        g = b.build()
        slow = O.sample_words(g, seed, [len(ev) - 1], 2 * grp, 2, 32, 3, fast=False)
        fast = O.sample_words(g, seed, [len(ev) - 1], 2 * grp, 2, 32, 3, fast=True)
        assert np.array_equal(slow, fast)

    check()


def _forced_tables():
    full = np.zeros(K, dtype=np.uint64)                                     # every group draws 32 ones
    stair = np.array([0] * 20 + [1 << 63] * 12, dtype=np.uint64)            # 20 or 32 ones per group
    return {"all32": full, "20or32": stair}


@pytest.mark.parametrize("name", ["all32", "20or32"])
def test_forced_counts_reach_the_third_and_fourth_position_call(name):
    """With the count table overridden (test hook) every group needs 20 to 40+ position fields, i.e. calls 0x401, 0x402 and
    often 0x403: the two oracle forms agree there too."""
    tab = _forced_tables()[name]
    a = O.sparse_words_with_table(2024, 7, tab, 128 * 5, 512, stream=True)
    b = O.sparse_words_with_table(2024, 7, tab, 128 * 5, 512, stream=False)
    assert np.array_equal(a, b)
    per_group = np.unpackbits(a.view(np.uint8)).reshape(-1, 128).sum(axis=1)
    assert set(per_group.tolist()) <= ({32} if name == "all32" else {20, 32}) and per_group.max() == 32


def test_law_of_the_sparse_bits():
    """Frequencies within 4 sigma of thr/2^32; ones per group distributed as Binomial(128, p); positions uniform."""
    b = G.GraphBuilder()
    ps = [1e-4, 3e-4, 1e-3, 2.0 ** -8 * 0.999, 0.0078, 0.031]
    for p in ps:
        b.root(b.gate(G.NULL, [b.basic_event(p)]))
    g = b.build()
    n = 1 << 24
    t = O.mc_packed(g, 99, 12345, n, nthreads=CORES, spec=3).astype(np.float64)
    q = np.array([O.quantised_p(p) for p in ps])
    assert np.all(np.abs(t / n - q) <= 4.0 * np.sqrt(q * (1 - q) / n))
    n_grp = 1 << 17
    for idx, kmax in ((4, 5), (5, 11)):     # about 1 and about 4 expected ones per group
        w = O.sample_words(g, 7, [idx], 0, 2 * n_grp, 32, 3, fast=True)[0]
        bits = np.unpackbits(w.view(np.uint8), bitorder="little").reshape(n_grp, 128)
        counts = bits.sum(axis=1)
        for k in range(kmax):
            pk = comb(128, k) * q[idx] ** k * (1 - q[idx]) ** (128 - k)
            assert abs((counts == k).sum() - n_grp * pk) <= 4.5 * np.sqrt(n_grp * pk * (1 - pk))
        per_pos = bits.sum(axis=0).astype(np.float64)
        assert np.all(np.abs(per_pos - n_grp * q[idx]) <= 4.5 * np.sqrt(n_grp * q[idx]))
        # neighbouring samples of a group are uncorrelated (the joint law is the product law)
        both = (bits[:, :-1] & bits[:, 1:]).sum()
        assert abs(both - n_grp * 127 * q[idx] ** 2) <= 4.5 * np.sqrt(n_grp * 127 * q[idx] ** 2)


@pytest.mark.parametrize("seed", range(4))
def test_mc_within_4_sigma_of_exact_enumeration(seed):
    g = random_graph(seed + 900, n_basic=10, n_gates=14, n_roots=2, extreme_p=False, consts=False, p_lo=0.001, p_hi=0.031)
    exact = O.exact(g, True)
    n = 1 << 20
    t = O.mc_packed(g, seed, 0, n, True, nthreads=CORES, spec=3).astype(np.float64)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / n)
    assert np.all(np.abs(t / n - exact) <= 4.0 * sigma + 3.0 / n)   # + 3 counts: gates with n p << 1 are Poisson, not normal


def test_events_are_independent_of_each_other():
    """AND of two or three sparse events fires with the product of their probabilities."""
    b = G.GraphBuilder()
    ps = [0.02, 0.03, 0.0078, 0.001, 0.031, 0.015]
    e = [b.basic_event(p) for p in ps]
    pairs = [(0, 1), (0, 2), (1, 4), (2, 3), (4, 5), (0, 5)]
    for i, j in pairs:
        b.root(b.gate(G.AND, [e[i], e[j]]))
    b.root(b.gate(G.AND, [e[0], e[1], e[4]]))
    n = 1 << 24
    t = O.mc_packed(b.build(), 4242, 0, n, nthreads=CORES, spec=3).astype(np.float64)
    q = [O.quantised_p(p) for p in ps]
    ex = np.array([q[i] * q[j] for i, j in pairs] + [q[0] * q[1] * q[4]])
    assert np.all(np.abs(t / n - ex) <= 4.0 * np.sqrt(ex * (1 - ex) / n))


def test_scalar_packed_and_lowered_agree_under_spec3():
    g = random_graph(31, n_basic=14, n_gates=30, n_roots=3, p_lo=1e-3, p_hi=0.02)
    want = O.mc_scalar(g, 5, 777, 3000, True, spec=3)
    assert np.array_equal(O.mc_packed(g, 5, 777, 3000, True, nthreads=3, spec=3), want)
    assert np.array_equal(O.mc_packed(g, 5, 777, 1000, True, spec=3) + O.mc_packed(g, 5, 1777, 2000, True, spec=3), want)
    with Plan(g, tally_all_gates=True, sampler=3) as plan:     # host lowering only: no GPU needed
        assert plan.spec == 3
        assert np.array_equal(eval_lowered(plan.view(), g, 5, 777, 3000, spec=3), want)


# ---------------------------------------------------------------------------------------------------------------
# GPU parity
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
def test_gpu_sampler_words_every_live_event(bits):
    g = synth.config_graph("c2", 0.05)
    with Plan(g, threshold_bits=bits, sampler=3) as plan:
        v = plan.view()
        got = plan.sample_words(seed=(1 << 63) + 12345, first_word=64 * 1000, n_words=128)
    assert np.array_equal(got, O.sample_words(g, (1 << 63) + 12345, v["event_orig"], 64 * 1000, 128, bits, 3, fast=True))


@pytest.mark.gpu
def test_gpu_sparse_boundary_thresholds_against_the_definition():
    b = G.GraphBuilder()
    ps = [1e-9, 3e-5, 2.0 ** -8, P_SPARSE - 2.0 ** -31, P_SPARSE, 0.0039, 0.1, 0.5, 2.0 ** -31, 1e-4, 0.0078, 0.02, 0.031]
    b.root(b.gate(G.OR, [b.basic_event(p) for p in ps]))
    g = b.build()
    with Plan(g, sampler=3) as plan:
        v = plan.view()
        got = plan.sample_words(seed=77, first_word=64 * 3, n_words=128)
    assert np.array_equal(got, O.sample_words(g, 77, v["event_orig"], 64 * 3, 128, 32, 3, fast=False))


@pytest.mark.gpu
@pytest.mark.parametrize("ev,grp,ones", EXT_GROUPS)
def test_gpu_groups_that_need_a_second_call(ev, grp, ones):
    g = _ext_graph()
    w0 = (2 * grp) // 128 * 128
    with Plan(g, sampler=3) as plan:
        got = plan.sample_words(seed=2024, first_word=w0, n_words=128)
    want = O.sample_words(g, 2024, range(4), w0, 128, 32, 3, fast=True)
    assert np.array_equal(got, want)
    assert _popcount(got[ev, 2 * grp - w0: 2 * grp - w0 + 2]) == ones


@pytest.mark.gpu
def test_gpu_long_sparse_stream_matches_oracle():
    """2^20 groups of four events at p = 0.0078: every count the table can give up to ~9 and many duplicate fields."""
    g = _ext_graph()
    n_words = 1 << 21
    with Plan(g, sampler=3) as plan:
        got = plan.sample_words(seed=2024, first_word=0, n_words=n_words)
    assert np.array_equal(got, O.sample_words(g, 2024, range(4), 0, n_words, 32, 3, fast=True))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["all32", "20or32"])
def test_gpu_forced_counts_reach_the_third_and_fourth_position_call(name):
    """Kernel against oracle where every group consumes 20 to 40+ position fields (count table overridden by the test hook)."""
    tab = _forced_tables()[name]
    b = G.GraphBuilder()
    ev = [b.basic_event(p) for p in (0.5, 0.01, 0.0078, 0.2, 0.02)]
    b.root(b.gate(G.OR, ev))
    g = b.build()
    n_words = 128 * 64
    with Plan(g, sampler=3) as plan:
        orig = [int(x) for x in plan.view()["event_orig"]]
        before = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
        plan.debug_set_sparse_table(orig.index(2), tab)
        got = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
        with pytest.raises(Exception):
            plan.debug_set_sparse_table(orig.index(0), tab)          # p = 0.5 is not a sparse event
        plan.update_p([0.5, 0.01, 0.0078, 0.2, 0.02])                # restores the real table
        after = plan.sample_words(seed=2024, first_word=128 * 5, n_words=n_words)
    want = O.sample_words(g, 2024, orig, 128 * 5, n_words, 32, 3, fast=True)
    assert np.array_equal(before, want) and np.array_equal(after, want)
    want[orig.index(2)] = O.sparse_words_with_table(2024, 2, tab, 128 * 5, n_words, stream=True)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_long_stream_near_four_ones_per_group():
    """p = 0.031 (3.97 expected ones per group): long count-table walks, the second position call (more than 8 fields) is
    routine.  A third call needs more than 24 fields, about 1e-11 per group at a real probability: the forced-count tests above
    cover it."""
    b = G.GraphBuilder()
    b.root(b.gate(G.OR, [b.basic_event(p) for p in (0.031, 0.031, 0.025, 0.016)]))
    g = b.build()
    n_words = 1 << 20
    with Plan(g, sampler=3) as plan:
        got = plan.sample_words(seed=99, first_word=128 * 7, n_words=n_words)
    want = O.sample_words(g, 99, range(4), 128 * 7, n_words, 32, 3, fast=True)
    assert np.array_equal(got, want)
    per_group = np.unpackbits(want[0].view(np.uint8)).reshape(-1, 128).sum(axis=1)
    assert per_group.max() >= 16 and (per_group > 8).mean() > 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(6))
def test_gpu_random_graphs_all_connectives(seed):
    g = random_graph(300 + seed, n_basic=16, n_gates=48, max_fanin=9, n_roots=4, p_lo=2e-3, p_hi=0.05)
    with Plan(g, tally_all_gates=True, sampler=3) as plan:
        got, _ = plan.run(20_000, seed=seed, sample_begin=1000 * seed + 7)
    assert np.array_equal(got, O.mc_scalar(g, seed, 1000 * seed + 7, 20_000, True, spec=3))


@pytest.mark.gpu
@pytest.mark.parametrize("name,scale,n", [("c2", 0.1, 1 << 15), ("c3", 0.05, 1 << 15), ("c4", 0.02, 1 << 14), ("c2", 1.0, 1 << 13)])
def test_gpu_configs(name, scale, n):
    g = synth.config_graph(name, scale)
    begin = (1 << 40) + 4096 * 9 + 13
    with Plan(g, tally_all_gates=True, sampler=3) as plan:
        got, rep = plan.run(n, seed=77, sample_begin=begin)
        halves = plan.run(n // 3, seed=77, sample_begin=begin)[0] + plan.run(n - n // 3, seed=77, sample_begin=begin + n // 3)[0]
    assert np.array_equal(got, O.mc_packed(g, 77, begin, n, True, nthreads=CORES, spec=3))
    assert np.array_equal(got, halves)


@pytest.mark.gpu
def test_gpu_update_p_moves_events_across_the_sparse_boundary():
    b = G.GraphBuilder()
    e = [b.basic_event(p) for p in (0.001, 0.05, 0.004, 0.3)]
    b.root(b.gate(G.OR, e[:2]))
    b.root(b.gate(G.ATLEAST, e, 2))
    g = b.build()
    n = 1 << 16
    with Plan(g, tally_all_gates=True, sampler=3) as plan:
        first, _ = plan.run(n, seed=11)
        plan.update_p([0.02, 0.0005, 0.004, 0.006])
        second, _ = plan.run(n, seed=11)
    g2 = G.from_lists([0.02, 0.0005, 0.004, 0.006], [(G.OR, [0, 1], 0, 0), (G.ATLEAST, [0, 1, 2, 3], 2, 0)], [4, 5])
    assert np.array_equal(first, O.mc_packed(g, 11, 0, n, True, nthreads=CORES, spec=3))
    assert np.array_equal(second, O.mc_packed(g2, 11, 0, n, True, nthreads=CORES, spec=3))
