"""Host logic: lowering pass, validation rules of the reference, C-ABI surface, shard helper.  CPU only --
nothing here computes on a GPU and nothing routes the product through the oracle: the oracle only checks
the lowered arrays, which are evaluated by a numpy evaluator that lives in tests/helpers.py."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from helpers import G, O, eval_lowered, random_graph, ROOT
from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan, Settings, SettingsError, estimate, shard_range


def test_library_exports_every_declared_symbol():
    """libcanopy.so exports exactly what include/canopy.h declares -- no test hook; libcanopy_testing.so adds exactly what
    include/canopy-testing.h declares."""
    pub = open(os.path.join(ROOT, "include", "canopy.h")).read()
    tst = open(os.path.join(ROOT, "include", "canopy-testing.h")).read()
    assert "debug" not in pub     # test hooks live in canopy-testing.h only
    rx = r"CANOPY_API\s+[\w\s\*]+?\b(canopy_\w+)\s*\("
    declared, hooks = set(re.findall(rx, pub)), set(re.findall(rx, tst))
    assert declared == set(_lib.SYMBOLS), (declared ^ set(_lib.SYMBOLS))
    assert hooks == set(_lib.TESTING_SYMBOLS), (hooks ^ set(_lib.TESTING_SYMBOLS))
    product = C.CDLL(os.path.join(os.path.dirname(_lib.TESTING_LIB_PATH), "libcanopy.so"))
    testing = C.CDLL(_lib.TESTING_LIB_PATH)
    for name in declared:
        assert hasattr(product, name) and hasattr(testing, name), name
    for name in hooks:
        assert hasattr(testing, name) and not hasattr(product, name), name
    assert _lib.lib().canopy_spec_version() == 4
    assert _lib.lib().canopy_status_string(3) == b"cycle error"


@pytest.mark.parametrize("seed", range(60))
def test_lowered_graph_equals_oracle_on_original(seed):
    g = random_graph(seed, n_basic=10, n_gates=25, n_roots=3)
    for tally_all in (False, True):
        with Plan(g, tally_all_gates=tally_all) as plan:
            got = eval_lowered(plan.view(), g, seed + 7, 100, 700)
        assert np.array_equal(got, O.mc_scalar(g, seed + 7, 100, 700, tally_all))


def test_lowered_structure_invariants():
    g = synth.generate(400, 220, 7, 24, seed=5)
    with Plan(g, tally_all_gates=True) as plan:
        v, st = plan.view(), plan.stats
    n_ev = v["n_events"]
    assert st["n_layers"] == v["n_layers"] and v["layer_off"][0] == 0 and v["layer_off"][-1] == v["n_gates"]
    assert list(v["event_orig"]) == sorted(v["event_orig"])
    level = np.zeros(n_ev + v["n_gates"], dtype=np.int64)
    for l in range(v["n_layers"]):
        fan = []
        for gi in range(v["layer_off"][l], v["layer_off"][l + 1]):
            kids = v["gate_child"][v["gate_child_off"][gi]:v["gate_child_off"][gi + 1]] & 0x7FFFFFFF
            assert len(kids) >= 1 and kids.max() < n_ev + v["layer_off"][l]          # only earlier layers
            assert level[kids].max() == l                                           # ASAP level: one child is right below
            level[n_ev + gi] = l + 1
            fan.append(len(kids))
        assert fan == sorted(fan, reverse=True)                                      # big gates first in a layer
    stored = (v["gate_flags"] & 2) != 0
    used = np.zeros(v["n_gates"], dtype=bool)
    kid_nodes = v["gate_child"] & 0x7FFFFFFF
    used[kid_nodes[kid_nodes >= n_ev] - n_ev] = True
    assert np.array_equal(stored, used)
    assert st["n_gates_stored"] == int(stored.sum()) and st["n_edges_live"] == len(v["gate_child"])


def test_dead_code_and_constant_folding():
    b = G.GraphBuilder()
    e = [b.basic_event(p) for p in (0.1, 0.2, 0.0, 1.0, 0.3)]
    dead = b.gate(G.OR, [e[0], e[4]])                       # never referenced, not a root
    a = b.gate(G.AND, [e[0], e[3], G.GraphBuilder.constant(True)])   # e3 is certain -> alias of e0
    z = b.gate(G.AND, [e[1], e[2]])                         # e2 impossible -> constant false
    k = b.gate(G.ATLEAST, [e[0], e[1], e[3], G.GraphBuilder.constant(True)], min_number=2)  # two trues -> always true
    n = b.gate(G.NOT, [a])
    top = b.gate(G.OR, [n, z])
    for h in (top, z, k, a):
        b.root(h)
    with Plan(b.build()) as plan:
        v, st = plan.view(), plan.stats
        assert st["n_basic_live"] == 1 and list(v["event_orig"]) == [0]     # only e0 is ever sampled
        assert list(v["tally_kind"]) == [3, 0, 1, 2]                         # top = !e0, z = false, k = true, a = e0
        assert st["n_gates_live"] == 1                                       # one buffer gate carries e0's tally
        got = eval_lowered(v, b.build(), 3, 0, 2000)
    assert np.array_equal(got, O.mc_scalar(b.build(), 3, 0, 2000))
    assert dead is not None


BAD = [
    ("and with one arg", (G.AND, [0]), _lib.E_VALIDITY, "2 or more"),
    ("not with two args", (G.NOT, [0, 1]), _lib.E_VALIDITY, "only one"),
    ("xor with three args", (G.XOR, [0, 1, 2]), _lib.E_VALIDITY, "exactly 2"),
    ("atleast min < 2", (G.ATLEAST, [0, 1, 2], 1), _lib.E_VALIDITY, "less than 2"),
    ("atleast n <= min", (G.ATLEAST, [0, 1], 2), _lib.E_VALIDITY, "more arguments"),
    ("cardinality max > n", (G.CARDINALITY, [0, 1], 0, 3), _lib.E_VALIDITY, "max number"),
    ("cardinality min > max", (G.CARDINALITY, [0, 1, 2], 2, 1), _lib.E_VALIDITY, "greater than max"),
    ("duplicate arg", (G.OR, [0, 0 | G.COMPLEMENT]), _lib.E_VALIDITY, "Duplicate"),
    ("complement under not", (G.NOT, [0 | G.COMPLEMENT]), _lib.E_INVALID, "complement"),
    ("constant under not", (G.NOT, [G.CONST_TRUE]), _lib.E_INVALID, "constant"),
    ("child out of range", (G.OR, [0, 99]), _lib.E_INVALID, "out of range"),
    ("unknown connective", (11, [0, 1]), _lib.E_INVALID, "connective"),
]


@pytest.mark.parametrize("name,gate,status,msg", BAD, ids=[b[0] for b in BAD])
def test_reference_validation_rules(name, gate, status, msg):
    """Arity / min / max / nesting rules of reference src/mef/openpsa/event/event.cpp:109-204."""
    g = G.from_lists([0.1, 0.2, 0.3], [gate], [3])
    with pytest.raises(_lib.CanopyError) as ei:
        Plan(g)
    assert ei.value.status == status and msg in ei.value.message


def test_cycle_and_domain_errors():
    g = G.from_lists([0.1], [(G.OR, [0, 2]), (G.AND, [0, 1])], [1])     # gate0 -> gate1 -> gate0
    with pytest.raises(_lib.CanopyError) as ei:
        Plan(g)
    assert ei.value.status == _lib.E_CYCLE
    for p in (-0.1, 1.5, float("nan")):
        with pytest.raises(_lib.CanopyError) as ei:
            Plan(G.from_lists([p, 0.5], [(G.OR, [0, 1])], [2]))
        assert ei.value.status == _lib.E_DOMAIN
    with pytest.raises(_lib.CanopyError) as ei:
        Plan(G.from_lists([0.5, 0.5], [(G.OR, [0, 1])], [7]))
    assert ei.value.status == _lib.E_INVALID


def test_empty_and_degenerate_graphs():
    with Plan(G.from_lists([], [], [])) as plan:
        assert plan.n_tallies == 0 and plan.stats["n_gates_live"] == 0
    with Plan(G.from_lists([0.25], [], [0])) as plan:                     # a basic event as the only root
        v = plan.view()
        assert v["n_gates"] == 1 and list(v["tally_kind"]) == [2]
        assert np.array_equal(eval_lowered(v, plan.graph, 1, 0, 640), O.mc_scalar(plan.graph, 1, 0, 640))


def test_wide_thresholds_are_exported():
    g = G.from_lists([2.0 ** -40, 0.5], [(G.OR, [0, 1])], [2])
    with Plan(g, threshold_bits=32) as p32, Plan(g, threshold_bits=64) as p64:
        assert p32.stats["n_basic_live"] == 1                 # 2^-40 is below 32-bit resolution: folded to false
        v = p64.view()
        assert p64.stats["n_basic_live"] == 2 and int(v["event_thr64"][0]) == 1 << 24


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_range_partitions_exactly(world):
    for begin, n in ((0, 1 << 20), (77, 100_003), (4096 * 5, 4096), (10, 5), (0, 0)):
        cuts = [shard_range(begin, n, r, world) for r in range(world)]
        assert cuts[0][0] == begin and sum(c[1] for c in cuts) == n
        for (b0, n0), (b1, _) in zip(cuts, cuts[1:]):
            assert b0 + n0 == b1
        for b, m in cuts[1:]:
            assert m == 0 or b % 4096 == 0 or b == begin      # interior cuts fall on slab boundaries


def test_settings_mirror_and_estimate():
    with pytest.raises(SettingsError):
        Settings(num_trials=0).validate()
    with pytest.raises(SettingsError):
        Settings(seed=-1).validate()
    p, s, lo, hi = estimate(250, 1000)
    assert p == 0.25 and abs(s - (0.25 * 0.75 / 1000) ** 0.5) < 1e-15 and lo < p < hi
    with pytest.raises(_lib.CanopyError):
        estimate(5, 0)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with Plan(random_graph(1)) as plan:
        with pytest.raises(_lib.CanopyError) as ei:
            plan.run(1000)
    assert ei.value.status == _lib.E_CUDA


def test_plan_cache_hits_are_verified_by_content():
    """canopy_plan_create keeps the last lowered forms keyed by a hash of the descriptor; a hit must be the same plan as a
    fresh lowering, and any change of an array or an option must miss."""
    from canopy_b200.engine import plan_cache_clear
    from canopy_b200 import synth
    plan_cache_clear()
    g = synth.generate(n_basic=400, n_gates=250, n_layers=7, max_fanin=9, seed=5)
    with Plan(g, True) as a, Plan(g, True) as b, Plan(g, True, plan_cache=False) as c:
        assert (a.stats["plan_cache_hit"], b.stats["plan_cache_hit"], c.stats["plan_cache_hit"]) == (0, 1, 0)
        fp = a.fingerprint()
        assert fp == b.fingerprint() == c.fingerprint()
        va, vb = a.view(), b.view()
        assert all(np.array_equal(va[k], vb[k]) for k in va)
        b.update_p(np.clip(g.p * 1.5, 0, 0.9))                  # a plan owns its copy: the cached form stays untouched
        assert b.fingerprint() != a.fingerprint()
    with Plan(g, True) as d:
        assert d.stats["plan_cache_hit"] == 1 and d.fingerprint() == fp
    with Plan(g, False) as e, Plan(g, True, threshold_bits=64) as f, Plan(g, True, sampler=2) as h:
        assert (e.stats["plan_cache_hit"], f.stats["plan_cache_hit"], h.stats["plan_cache_hit"]) == (0, 0, 0)
    g2 = G.Graph(p=g.p.copy(), op=g.op, kmin=g.kmin, kmax=g.kmax, child_off=g.child_off, child=g.child.copy(), roots=g.roots)
    g2.child[7] ^= 0x80000000
    with Plan(g2, True) as x:
        assert x.stats["plan_cache_hit"] == 0
    for i in range(6):                                          # more graphs than the cache holds: the oldest entry goes
        with Plan(synth.generate(50 + i, 30, 4, 5, seed=i)) as y:
            assert y.stats["plan_cache_hit"] == 0
    with Plan(g, True) as z:
        assert z.stats["plan_cache_hit"] == 0
    plan_cache_clear()


def test_device_child_stream_layout():
    """The child stream the gate walk reads (csrc/lowering.cpp): per layer the runs of the walk's warps, each a whole number of
    four-code groups (the pad repeats the run's first row), gate arguments in canonical order inside their run."""
    import ctypes as C
    from canopy_b200 import _lib, synth
    from canopy_b200.engine import Plan
    for g in (synth.generate(n_basic=300, n_gates=200, n_layers=7, max_fanin=12, seed=5), random_graph(7, n_basic=30, n_gates=80, max_fanin=9)):
        with Plan(g, tally_all_gates=True) as plan:
            v = plan.view()
            st = _lib.CPlanStreams()
            _lib.check(_lib.lib().canopy_plan_debug_streams(plan._h, C.byref(st)))
            W, nl, ng = st.warps, st.n_layers, int(v["n_gates"]) if "n_gates" in v else len(v["gate_op"])
            dev = np.ctypeslib.as_array(st.dev_child, shape=(st.n_dev_child,)).copy()
            off = np.ctypeslib.as_array(st.gate_child_off, shape=(ng,)).copy()
            split = np.ctypeslib.as_array(st.split, shape=(nl, W + 1)).copy()
            sc = np.ctypeslib.as_array(st.split_child, shape=(nl, W + 1)).copy()
            canon, coff = v["gate_child"], v["gate_child_off"]
            n_nodes = len(v["event_orig"]) + ng
            assert st.n_dev_child % 4 == 0 and np.all((dev & 0x7FFFFFFF) < max(n_nodes, 1))
            for g_i in range(ng):   # every gate's arguments, contiguous and in canonical order
                n = int(coff[g_i + 1] - coff[g_i])
                assert np.array_equal(dev[off[g_i]: off[g_i] + n], canon[coff[g_i]: coff[g_i + 1]])
            for l in range(nl):
                assert np.all(np.diff(split[l]) >= 0) and split[l, 0] == v["layer_off"][l]
                for w in range(W):
                    a, b = int(sc[l, w]), int(sc[l, w + 1])
                    assert a % 4 == 0 and (b - a) % 4 == 0
                    gates = range(int(split[l, w]), int(split[l, w + 1]))
                    real = sum(int(coff[x + 1] - coff[x]) for x in gates)
                    assert 0 <= (b - a) - real < 4
                    if gates:
                        assert off[gates[0]] == a
                        assert np.all(dev[a + real: b] == (dev[a] & 0x7FFFFFFF))   # the pad: a row that exists, plain
                    else:
                        assert a == b
