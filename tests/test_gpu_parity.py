"""GPU parity tests proper: the CUDA path through the C ABI against the oracle, bit-exact.
Sizes are chosen so the oracle (multi-threaded packed port, itself checked against the scalar oracle on CPU)
finishes in seconds; BASELINE-sized runs are covered by size-independent properties in test_gpu_properties.py."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import G, O, ROOT, random_graph
from canopy_b200 import synth
from canopy_b200.engine import Plan, load_mef

pytestmark = pytest.mark.gpu
CORES = len(os.sched_getaffinity(0))


def _golden():
    with open(os.path.join(ROOT, "tests", "golden", "golden_tallies.json")) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("case", _golden(), ids=lambda c: c["name"])
def test_golden_fixtures(case):
    g = case["graph"]
    graph = G.Graph(p=g["p"], op=g["op"], kmin=g["kmin"], kmax=g["kmax"], child_off=g["child_off"], child=g["child"],
                    roots=g["roots"])
    with Plan(graph, case["tally_all"], case["bits"], sampler=case["spec"]) as plan:
        assert plan.spec == case["spec"]
        got, _ = plan.run(case["n"], seed=case["seed"], sample_begin=case["begin"])
        ev = [int(np.nonzero(plan.view()["event_orig"] == e)[0][0]) for e in case["word_events"]
              if e in plan.view()["event_orig"]]
        w0 = case["first_word"] // 64 * 64
        words = plan.sample_words(case["seed"], w0, 128)[ev][:, case["first_word"] - w0: case["first_word"] - w0 + 2]
    assert np.array_equal(got, np.array(case["tallies"], dtype=np.uint64))
    if len(ev) == len(case["word_events"]):
        assert np.array_equal(words, np.array(case["words"], dtype=np.uint64).reshape(words.shape))


def test_c1_mef_example_2pow20():
    """BASELINE.json config 1: the small AND/OR MEF tree, 2^20 samples, GPU vs the CPU path."""
    g = load_mef(os.path.join(ROOT, "examples", "c1_small_ft.xml"))
    n = 1 << 20
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(n, seed=2024)
    assert np.array_equal(got, O.mc_packed(g, 2024, 0, n, True, nthreads=CORES))


@pytest.mark.parametrize("name,scale,n", [("c2", 0.1, 1 << 15), ("c3", 0.05, 1 << 15), ("c4", 0.02, 1 << 14)])
def test_scaled_configs(name, scale, n):
    g = synth.config_graph(name, scale)
    tally_all = name == "c3"
    with Plan(g, tally_all) as plan:
        got, rep = plan.run(n, seed=77, sample_begin=4096 * 9 + 13)
    assert np.array_equal(got, O.mc_packed(g, 77, 4096 * 9 + 13, n, tally_all, nthreads=CORES))
    assert rep.gate_launches >= 1 and rep.sampler_launches >= 1


@pytest.mark.parametrize("name,n", [("c3", 8192 * 2 + 4096 + 77), ("c4", 8192 + 1234)])
def test_c3_c4_full_graphs_short_run_per_gate(name, n):
    """BASELINE configs 3 and 4 at their FULL graph size (20.2k gates with 200 sequence roots; 100k events with fan-in up
    to 245, 36 layers), every gate tallied, over a ragged sample range that straddles slabs: bit-exact against the oracle."""
    g = synth.config_graph(name)
    begin = (1 << 40) + 8192 * 5 + 4000
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(n, seed=0xC0DE, sample_begin=begin)
    assert np.array_equal(got, O.mc_packed(g, 0xC0DE, begin, n, True, nthreads=CORES))


def test_c2_full_graph_short_run():
    g = synth.config_graph("c2")
    n = 1 << 13
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(n, seed=5)
    assert np.array_equal(got, O.mc_packed(g, 5, 0, n, True, nthreads=CORES))


@pytest.mark.parametrize("spec", [1, 2, 4])
@pytest.mark.parametrize("bits", [32, 64])
def test_sampler_words_every_live_event(spec, bits):
    g = synth.config_graph("c2", 0.05)
    with Plan(g, threshold_bits=bits, sampler=spec) as plan:
        v = plan.view()
        got = plan.sample_words(seed=(1 << 63) + 12345, first_word=64 * 1000, n_words=64)
    assert np.array_equal(got, O.sample_words(g, (1 << 63) + 12345, v["event_orig"], 64 * 1000, 64, bits, spec, fast=True))


@pytest.mark.parametrize("bits", [32, 64])
def test_sliced_sampler_threshold_bytes(bits):
    """Spec 2 decisions for thresholds whose top byte is 0x00 / 0xFF / mixed and whose tail is zero, against the
    oracle's full-variate definition (fast=False)."""
    b = G.GraphBuilder()
    ps = [1e-9, 3e-5, 2.0 ** -8, 2.0 ** -8 + 2.0 ** -30, 0.0039, 0.1, 0.5, 0.75, 1 - 2.0 ** -9, 0.999999, 2.0 ** -31]
    b.root(b.gate(G.OR, [b.basic_event(p) for p in ps]))
    g = b.build()
    with Plan(g, threshold_bits=bits, sampler=2) as plan:
        v = plan.view()
        got = plan.sample_words(seed=77, first_word=64 * 3, n_words=128)
    assert np.array_equal(got, O.sample_words(g, 77, v["event_orig"], 64 * 3, 128, bits, 2, fast=False))


@pytest.mark.parametrize("seed", range(12))
def test_random_graphs_all_connectives(seed):
    g = random_graph(100 + seed, n_basic=16, n_gates=48, max_fanin=9, n_roots=4)
    spec = (2, 4, 2, 1)[seed % 4]
    with Plan(g, tally_all_gates=True, sampler=spec) as plan:
        got, _ = plan.run(20_000, seed=seed, sample_begin=1000 * seed + 7)
    assert np.array_equal(got, O.mc_scalar(g, seed, 1000 * seed + 7, 20_000, True, spec=spec))


@pytest.mark.parametrize("spec", [1, 2, 4])
@pytest.mark.parametrize("bits", [32, 64])
def test_rare_events_and_threshold_width(bits, spec):
    """p down to 1e-9: 32-bit thresholds quantise to multiples of 2^-32, 64-bit ones resolve them."""
    b = G.GraphBuilder()
    e = [b.basic_event(p) for p in (1e-9, 3e-7, 1e-4, 0.02, 0.5, 2.0 ** -33)]
    b.root(b.gate(G.OR, e))
    b.root(b.gate(G.AND, [e[3], e[4]]))
    g = b.build()
    n = 1 << 17
    with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=spec) as plan:
        got, _ = plan.run(n, seed=31337)
    assert np.array_equal(got, O.mc_packed(g, 31337, 0, n, True, bits=bits, nthreads=CORES, spec=spec))


@pytest.mark.parametrize("fanin,lo,hi", [(5, 2, 0), (40, 7, 0), (300, 30, 0), (300, 20, 40), (700, 95, 0), (700, 90, 100), (33, 0, 3)])
def test_high_fanin_count_gates(fanin, lo, hi):
    """k-of-n / cardinality across every bit-sliced counter width (3, 6, 9 and 16 planes)."""
    b = G.GraphBuilder()
    e = [b.basic_event(0.02 + 0.1 * ((i * 37) % 11) / 11) for i in range(fanin)]
    args = [(x, i % 13 == 0) for i, x in enumerate(e)]
    if hi:
        b.root(b.gate(G.CARDINALITY, args, lo, hi))
    else:
        b.root(b.gate(G.ATLEAST, args, lo))
    b.root(b.gate(G.OR, e[:fanin // 2] ))
    b.root(b.gate(G.AND, [(x, True) for x in e]))
    g = b.build()
    n = 3 * 4096
    with Plan(g) as plan:
        got, _ = plan.run(n, seed=fanin)
    want = O.mc_packed(g, fanin, 0, n, nthreads=CORES)
    assert np.array_equal(got, want)
    assert 0 < int(got[0]) < n          # the threshold sits inside the count distribution: the gate is exercised


_ORACLE_CACHE = {}   # the two engine modes of conftest.py both take the streamed path at these sizes: ask the oracle once


def _oracle_cached(name, g, seed, begin, n, bits):
    key = (name, seed, begin, n, bits)
    if key not in _ORACLE_CACHE:
        _ORACLE_CACHE[key] = O.mc_packed(g, seed, begin, n, True, bits=bits, nthreads=CORES)
    return _ORACLE_CACHE[key]


@pytest.mark.parametrize("name,bits", [("c4", 32), ("c4", 64), ("c3", 32)])
def test_full_graphs_pipelined_over_several_batches_per_gate(name, bits):
    """The benched configuration on the path that produces the bench number: FULL c4 (and c3) graph, every gate tallied,
    default batch size (two slabs per SM) and the default overlapped pipeline -- sampler of batch b+1 beside the walk of
    batch b on two value buffers, each buffer rewritten at least twice -- over a range that starts inside a slab and ends
    with a ragged tail, against the oracle (about 30 s of host time per case on 16 cores)."""
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    g = synth.config_graph(name)
    begin = (1 << 40) + 8192 * 3 + 4001
    n = 4 * (2 * sms) * 8192 + 8192 + 1234
    with Plan(g, tally_all_gates=True, threshold_bits=bits) as plan:
        got, rep = plan.run(n, seed=0xFACADE, sample_begin=begin)
    # the default batch is two slabs per SM, less when two value buffers of that size exceed half the free memory (c4: 293 of 296)
    assert rep.n_batches >= 5 and sms <= rep.batch_slabs <= 2 * sms and rep.sampler_launches == rep.n_batches
    want = _oracle_cached(name, g, 0xFACADE, begin, n, bits)
    assert want[:g.n_roots].sum() > 0
    assert np.array_equal(got, want)


@pytest.fixture
def row_copies():
    """include/canopy-testing.h: make every gate walk take gate_ring_kernel_t<false> (one bulk copy per row), the path the
    engine itself only chooses for value buffers beyond 2^31 rows or without a tensor-map encoder."""
    from canopy_b200 import _lib
    _lib.check(_lib.lib().canopy_debug_force_row_copies(1))
    yield
    _lib.check(_lib.lib().canopy_debug_force_row_copies(0))


@pytest.mark.parametrize("name,scale,n,batch", [("c2", 0.1, 1 << 15, 0), ("c3", 0.05, 1 << 15, 0), ("c4", 0.02, 1 << 14, 0),
                                                ("c4", 0.05, 8192 * 50 + 99, 8), ("c2", 1.0, 8192 * 40 + 5, 6)])
def test_walk_with_one_bulk_copy_per_row(row_copies, name, scale, n, batch):
    """The walk's fallback instantiation against the oracle: scaled configs (k-of-n, wide gates, fan-in up to 245), and two
    pipelined runs of 7 batches on two value buffers; then the same run on the default (gather4) path gives the same tallies."""
    from canopy_b200 import _lib
    from canopy_b200.engine import ENGINE_STREAMED
    g = synth.config_graph(name, scale)
    begin = 4096 * 9 + 13
    with Plan(g, tally_all_gates=True, engine=ENGINE_STREAMED) as plan:
        got, rep = plan.run(n, seed=77, sample_begin=begin, batch_slabs=batch)
        assert rep.gate_launches >= (7 if batch else 1)
        _lib.check(_lib.lib().canopy_debug_force_row_copies(0))
        again, _ = plan.run(n, seed=77, sample_begin=begin, batch_slabs=batch)
    assert np.array_equal(got, O.mc_packed(g, 77, begin, n, True, nthreads=CORES))
    assert np.array_equal(got, again)


@pytest.mark.parametrize("seed", range(6))
def test_walk_with_one_bulk_copy_per_row_all_connectives(row_copies, seed):
    from canopy_b200.engine import ENGINE_STREAMED
    g = random_graph(500 + seed, n_basic=16, n_gates=48, max_fanin=9, n_roots=4)
    with Plan(g, tally_all_gates=True, engine=ENGINE_STREAMED) as plan:
        got, _ = plan.run(20_000, seed=seed, sample_begin=1000 * seed + 7)
    assert np.array_equal(got, O.mc_scalar(g, seed, 1000 * seed + 7, 20_000, True))
