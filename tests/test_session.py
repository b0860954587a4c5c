"""Restart record and probability update: host-side logic, no GPU."""
import numpy as np
import pytest

from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan, Session


@pytest.fixture(scope="module")
def graph():
    return synth.generate(n_basic=60, n_gates=40, n_layers=5, max_fanin=6, seed=11)


def test_fingerprint_is_a_function_of_graph_options_and_probabilities(graph):
    with Plan(graph) as a, Plan(graph) as b, Plan(graph, tally_all_gates=True) as c, Plan(graph, sampler=1) as d:
        assert a.fingerprint() == b.fingerprint()
        assert len({a.fingerprint(), c.fingerprint(), d.fingerprint()}) == 3
        live = a.view()["event_orig"]
        p = graph.p.copy(); p[live[0]] *= 0.5
        b.update_p(p)
        assert b.fingerprint() != a.fingerprint()
        v = b.view()
        assert v["event_thr32"][0] == int(p[live[0]] * 2 ** 32) and np.array_equal(v["event_thr32"][1:], a.view()["event_thr32"][1:])
        b.update_p(graph.p)
        assert b.fingerprint() == a.fingerprint()


def test_update_p_refuses_what_would_change_the_lowering(graph):
    with Plan(graph) as plan:
        live = int(plan.view()["event_orig"][0])
        for bad, status in ((0.0, _lib.E_INVALID), (1.0, _lib.E_INVALID), (1e-12, _lib.E_INVALID), (1.5, _lib.E_DOMAIN), (float("nan"), _lib.E_DOMAIN)):
            p = graph.p.copy(); p[live] = bad
            with pytest.raises(_lib.CanopyError) as ei:
                plan.update_p(p)
            assert ei.value.status == status
        with pytest.raises(_lib.CanopyError):
            plan.update_p(graph.p[:-1])
        assert plan.fingerprint() == Plan(graph).fingerprint()  # a refused update leaves the plan untouched


def test_checkpoint_record_round_trip_and_validation(graph, tmp_path):
    with Plan(graph) as plan, Plan(graph, threshold_bits=64) as other:
        s = Session(plan, seed=7, sample_begin=4096)
        assert s.n_done == 0 and s.ck.spec_version == plan.spec and s.ck.n_tallies == plan.n_tallies
        s.tallies[0] = 5
        f = str(tmp_path / "ck.npz")
        s.save(f)
        r = Session.load(f, plan)
        assert (r.ck.seed, r.ck.sample_begin, r.ck.next_sample, r.ck.plan_fingerprint) == (7, 4096, 4096, plan.fingerprint())
        assert int(r.tallies[0]) == 5
        # a record cannot be resumed on a different plan; the check comes before any device work
        with pytest.raises(_lib.CanopyError) as ei:
            Session.load(f, other).advance(64)
        assert ei.value.status == _lib.E_INVALID and "another plan" in ei.value.message
        r.ck.magic = 0
        with pytest.raises(_lib.CanopyError) as ei:
            r.advance(64)
        assert "bad magic" in ei.value.message
