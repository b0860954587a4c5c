"""Parameter-uncertainty source: the six deviates of reference expr/random_deviate.h:25-264 drawn from
counter-based Philox streams, Expression::Sample()/Reset() semantics (expression.h:98-117) on a loaded MEF model, and
periodic-test (expr/exponential.h:105-245).  CPU only: the library's draws are compared with the independent Python
restatement (tests/deviate_restatement.py, tolerance 1e-12 relative: libm calls only) and with closed-form moments.
The outer loop itself runs on the device: tests/test_gpu_uncertainty.py."""
import math

import numpy as np
import pytest

import deviate_restatement as R
from canopy_b200 import _lib
from canopy_b200.engine import (DEV_BETA, DEV_CONSTANT, DEV_GAMMA, DEV_HISTOGRAM, DEV_LOGNORMAL, DEV_LOGNORMAL_EF, DEV_NORMAL,
                                DEV_UNIFORM, PSource, load_mef)

CASES = [
    ("uniform", DEV_UNIFORM, (0.1, 0.3)),
    ("normal", DEV_NORMAL, (0.5, 0.1)),
    ("lognormal", DEV_LOGNORMAL, (-3.0, 0.5)),
    ("lognormal_ef", DEV_LOGNORMAL_EF, (1e-3, 3.0, 0.95)),
    ("gamma", DEV_GAMMA, (2.0, 0.05)),
    ("gamma", DEV_GAMMA, (0.4, 0.05)),
    ("beta", DEV_BETA, (2.0, 5.0)),
    ("beta", DEV_BETA, (0.5, 0.7)),
]
HIST = ([0.0, 0.1, 0.3, 0.4], [1.0, 2.0, 1.0])


def closed_form(kind, prm):
    """(mean, variance) of each distribution."""
    if kind == "uniform":
        return (prm[0] + prm[1]) / 2, (prm[1] - prm[0]) ** 2 / 12
    if kind == "normal":
        return prm[0], prm[1] ** 2
    if kind == "lognormal":
        m = math.exp(prm[0] + prm[1] ** 2 / 2)
        return m, m * m * (math.exp(prm[1] ** 2) - 1)
    if kind == "lognormal_ef":
        s = math.log(prm[1]) / 1.6448536269514722
        return prm[0], prm[0] ** 2 * (math.exp(s * s) - 1)
    if kind == "gamma":
        return prm[0] * prm[1], prm[0] * prm[1] ** 2
    if kind == "beta":
        a, b = prm
        return a / (a + b), a * b / ((a + b) ** 2 * (a + b + 1))
    raise ValueError(kind)


def test_draws_equal_the_python_restatement():
    dists = [(k,) + prm for _, k, prm in CASES] + [(DEV_HISTOGRAM,) + HIST, (DEV_CONSTANT, 0.25)]
    ps = PSource.from_deviates(dists)
    assert ps.n_basic == len(dists) and ps.n_deviates == len(dists) - 1
    for seed, trial in [(0, 0), (7, 1), (0xC0FFEE, 12345), (2 ** 40 + 3, 2 ** 33 + 5)]:
        got = ps.sample(seed, trial)
        for i, (name, _, prm) in enumerate(CASES):
            want = min(1.0, max(0.0, R.draw(name, prm, seed, trial, i)))
            assert got[i] == pytest.approx(want, rel=1e-12, abs=1e-300), (name, prm, seed, trial)
        assert got[len(CASES)] == pytest.approx(R.draw("histogram", HIST, seed, trial, len(CASES)), rel=1e-12)
        assert got[-1] == 0.25


def test_moments_against_closed_forms():
    n = 40000
    dists = [(k,) + prm for _, k, prm in CASES] + [(DEV_HISTOGRAM,) + HIST]
    ps = PSource.from_deviates(dists)
    # unclamped moments are wanted: every case above stays inside [0, 1] except in far tails (normal(0.5, 0.1): 5 sigma)
    x = np.stack([ps.sample(99, t) for t in range(n)])
    mean = ps.mean()
    for i, (name, _, prm) in enumerate(CASES):
        m, var = closed_form(name, prm)
        assert mean[i] == pytest.approx(m, rel=1e-12)
        assert abs(x[:, i].mean() - m) < 5 * math.sqrt(var / n), (name, prm)
        # variance of the sample variance ~ (kurtosis - 1) var^2 / n; 0.1 relative is > 5 sigma for all cases here
        assert x[:, i].var() == pytest.approx(var, rel=0.1), (name, prm)
    b, w = HIST
    hm = sum(wi * (b[i] + b[i + 1]) / 2 for i, wi in enumerate(w)) / sum(w)
    assert mean[-1] == pytest.approx(hm, rel=1e-12) and abs(x[:, -1].mean() - hm) < 5 * 0.12 / math.sqrt(n)
    # bin masses of the histogram deviate
    for i, wi in enumerate(w):
        frac = np.mean((x[:, -1] >= b[i]) & (x[:, -1] < b[i + 1]))
        assert abs(frac - wi / sum(w)) < 5 * math.sqrt(0.25 / n)


def test_trials_are_independent_and_reproducible():
    ps = PSource.from_deviates([(DEV_UNIFORM, 0.0, 1.0), (DEV_UNIFORM, 0.0, 1.0)])
    a = np.stack([ps.sample(5, t) for t in range(20000)])
    assert np.array_equal(a[:100], np.stack([ps.sample(5, t) for t in range(100)]))
    assert not np.array_equal(a[:100], np.stack([ps.sample(6, t) for t in range(100)]))
    assert abs(np.corrcoef(a[:, 0], a[:, 1])[0, 1]) < 0.03       # between deviates
    assert abs(np.corrcoef(a[:-1, 0], a[1:, 0])[0, 1]) < 0.03    # between consecutive trials


def test_validation_follows_the_reference_rules():
    for d in [(DEV_UNIFORM, 0.3, 0.1), (DEV_NORMAL, 0.1, 0.0), (DEV_LOGNORMAL_EF, 1e-3, 0.9, 0.95), (DEV_LOGNORMAL_EF, 1e-3, 3, 1.0),
              (DEV_GAMMA, 0.0, 1.0), (DEV_BETA, 1.0, -1.0), (DEV_HISTOGRAM, [0, 0.2, 0.1], [1, 1]), (DEV_HISTOGRAM, [0, 0.1], [0.0]),
              (99, 0.0)]:
        with pytest.raises(_lib.CanopyError):
            PSource.from_deviates([d])


UNC_XML = """<opsa-mef>
 <define-fault-tree name="ft">
  <define-gate name="top"><or><basic-event name="a"/><basic-event name="b"/><basic-event name="c"/><basic-event name="d"/></or></define-gate>
  <define-basic-event name="a"><exponential><parameter name="lambda"/><system-mission-time/></exponential></define-basic-event>
  <define-basic-event name="b"><mul><parameter name="lambda"/><float value="100"/></mul></define-basic-event>
  <define-basic-event name="c"><beta-deviate><float value="2"/><float value="50"/></beta-deviate></define-basic-event>
  <define-basic-event name="d"><float value="0.01"/></define-basic-event>
 </define-fault-tree>
 <model-data>
  <define-parameter name="lambda"><lognormal-deviate><float value="1e-4"/><float value="3"/><float value="0.95"/></lognormal-deviate></define-parameter>
 </model-data>
</opsa-mef>"""


def test_model_source_shares_a_parameter_between_its_users():
    g = load_mef(text=UNC_XML, mission_time=1000.0)
    ps = g.meta["psource"]
    assert ps.n_basic == 4 and ps.n_deviates == 2
    # mean mode reproduces the loader's probabilities exactly (same operations in the same order)
    assert np.array_equal(ps.mean(), g.p)
    names = g.basic_names
    ia, ib, ic, id_ = (names.index(x) for x in "abcd")
    seen = set()
    for t in range(50):
        p = ps.sample(11, t)
        lam = p[ib] / 100.0                                   # Expression::Sample(): ONE draw of lambda per trial ...
        assert p[ia] == pytest.approx(1 - math.exp(-lam * 1000.0), rel=1e-12)   # ... seen by both events
        assert p[id_] == 0.01
        seen.add((p[ib], p[ic]))
    assert len(seen) == 50                                    # Reset(): a new draw every trial
    # the parameter's deviate is deviate 0 or 1 of the tape; either way the draw matches the restatement
    lam0 = ps.sample(11, 0)[ib] / 100.0
    assert any(lam0 == pytest.approx(R.draw("lognormal_ef", (1e-4, 3.0, 0.95), 11, 0, d), rel=1e-12) for d in (0, 1))


def test_ccf_and_event_tree_probabilities_flow_from_sampled_parameters():
    xml = """<opsa-mef>
     <define-fault-tree name="ft">
      <define-gate name="top"><and><basic-event name="p1"/><basic-event name="p2"/><basic-event name="p3"/></and></define-gate>
      <define-CCF-group name="pumps" model="beta-factor">
       <members><basic-event name="p1"/><basic-event name="p2"/><basic-event name="p3"/></members>
       <distribution><parameter name="q"/></distribution>
       <factor level="3"><parameter name="beta"/></factor>
      </define-CCF-group>
     </define-fault-tree>
     <model-data>
      <define-parameter name="q"><uniform-deviate><float value="0.05"/><float value="0.15"/></uniform-deviate></define-parameter>
      <define-parameter name="beta"><uniform-deviate><float value="0.1"/><float value="0.3"/></uniform-deviate></define-parameter>
     </model-data>
    </opsa-mef>"""
    g = load_mef(text=xml, ccf_analysis=True)
    ps = g.meta["psource"]
    assert np.array_equal(ps.mean(), g.p)
    names = g.basic_names
    ind = [i for i, n in enumerate(names) if n in ("[p1]", "[p2]", "[p3]")]
    com = names.index("[p1 p2 p3]")
    assert len(ind) == 3
    for t in range(20):
        p = ps.sample(3, t)
        q = p[ind[0]] + p[com]                                # (1 - beta) q + beta q
        beta = p[com] / q
        assert 0.05 <= q <= 0.15 and 0.1 <= beta <= 0.3
        assert p[ind[0]] == p[ind[1]] == p[ind[2]]            # one draw of q and beta for the whole group


def test_periodic_test_values():
    def p_of(args):
        body = "".join(f'<float value="{a!r}"/>' for a in args)
        xml = f"""<opsa-mef><define-fault-tree name="ft"><define-gate name="top"><or><basic-event name="a"/><basic-event name="b"/></or></define-gate>
          <define-basic-event name="a"><periodic-test>{body}</periodic-test></define-basic-event>
          <define-basic-event name="b"><float value="0.5"/></define-basic-event></define-fault-tree></opsa-mef>"""
        g = load_mef(text=xml)
        return g.p[g.basic_names.index("a")]
    lam, tau, theta = 1e-3, 100.0, 30.0
    assert p_of([lam, tau, theta, 20.0]) == pytest.approx(1 - math.exp(-lam * 20.0), rel=1e-12)      # before the first test
    assert p_of([lam, tau, theta, 30.0 + 2 * 100.0 + 17.0]) == pytest.approx(1 - math.exp(-lam * 17.0), rel=1e-9)
    # five arguments: closed form in the library vs direct integration of the three-state model
    for mu, time in [(0.05, 20.0), (0.05, 130.0), (0.002, 477.0), (1e-3, 250.0)]:
        assert p_of([lam, mu, tau, theta, time]) == pytest.approx(R.periodic_test5_ode(lam, mu, tau, theta, time), rel=1e-6)
    # instant repair is the limit of a fast repair
    assert p_of([lam, 50.0, tau, theta, 477.0]) == pytest.approx(p_of([lam, tau, theta, 477.0]), rel=1e-2)
    for bad in ([lam, -1.0, theta, 10.0], [lam, tau, theta], [lam] * 11):
        with pytest.raises(_lib.CanopyError):
            p_of(bad)
