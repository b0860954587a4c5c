"""Parameter-uncertainty outer loop on the device (canopy_uncertainty_run): every epistemic trial draws all
deviates (counter-based), re-quantises the thresholds of one lowered plan and reruns the inner Monte Carlo.

Checks: (1) bit-exact parity of every trial's tallies with the scalar oracle run on a graph that carries the probabilities
of that trial (drawn through canopy_psource_sample, itself pinned by tests/test_uncertainty.py against the Python
restatement); (2) the outer-loop mean of p_hat against the mean-value quantification within 3 sigma (independent deviates
under an OR gate, where E[f(p)] = f(E[p])); (3) quantiles, bins and the restored plan.
Reference anchors: Settings::uncertainty_analysis / num_quantiles / num_bins settings.h:154-175,281-295;
Expression::Sample()/Reset() expression.h:98-117; expr/random_deviate.h:25-264."""
import math

import numpy as np
import pytest

from helpers import G, O
from canopy_b200 import synth
from canopy_b200.engine import (DEV_BETA, DEV_CONSTANT, DEV_GAMMA, DEV_HISTOGRAM, DEV_LOGNORMAL, DEV_LOGNORMAL_EF, DEV_UNIFORM,
                                Plan, PSource, Settings, load_mef, uncertainty)
from test_uncertainty import UNC_XML

pytestmark = pytest.mark.gpu


def with_p(g, p):
    return G.Graph(p=p, op=g.op, kmin=g.kmin, kmax=g.kmax, child_off=g.child_off, child=g.child, roots=g.roots)


@pytest.mark.parametrize("bits", [32, 64])
def test_every_trial_is_bit_exact_against_the_oracle(bits):
    g = synth.generate(n_basic=200, n_gates=120, n_layers=6, max_fanin=8, seed=77)
    kinds = [DEV_LOGNORMAL_EF, DEV_GAMMA, DEV_BETA, DEV_UNIFORM, DEV_CONSTANT]
    dists = []
    for i, p in enumerate(g.p):
        k = kinds[i % len(kinds)]
        dists.append({DEV_LOGNORMAL_EF: (k, p, 3.0, 0.95), DEV_GAMMA: (k, 2.0, p / 2.0), DEV_BETA: (k, 2.0, 2.0 / p - 2.0),
                      DEV_UNIFORM: (k, p / 2, min(1.0, 3 * p / 2)), DEV_CONSTANT: (k, p)}[k])
    ps = PSource.from_deviates(dists)
    assert np.allclose(ps.mean(), g.p, rtol=1e-12)
    n, n_outer, seed, dseed, begin = 2 * 8192 + 321, 5, 17, 4242, 4096 * 3
    st = Settings(num_trials=n, seed=seed, sample_begin=begin, tally_all_gates=True, threshold_bits=bits)
    u = uncertainty(g, st, ps, n_outer, deviate_seed=dseed)
    assert u["p"].shape == (n_outer, g.n_tallies(True))
    lo, hi = 2.0 ** -bits, 1.0 - 2.0 ** -53
    for t in range(n_outer):
        p = np.clip(ps.sample(dseed, t), lo, hi)
        want = O.mc_scalar(with_p(g, p), seed, begin + t * n, n, True, bits=bits)
        assert np.array_equal(np.rint(u["p"][t] * n).astype(np.uint64), want), t
        assert np.array_equal(u["p"][t], want.astype(np.float64) / n)


def test_outer_mean_matches_the_mean_value_run_within_3_sigma():
    # top = OR of 6 events with independent deviates: E[1 - prod(1 - p_i)] = 1 - prod(1 - E p_i)
    b = G.GraphBuilder()
    dists = [(DEV_LOGNORMAL_EF, 0.02, 3.0, 0.95), (DEV_GAMMA, 2.0, 0.01), (DEV_BETA, 2.0, 60.0), (DEV_UNIFORM, 0.01, 0.05),
             (DEV_LOGNORMAL, math.log(0.01), 0.4), (DEV_HISTOGRAM, [0.0, 0.02, 0.06], [3.0, 1.0])]
    ps = PSource.from_deviates(dists)
    ev = [b.basic_event(p) for p in ps.mean()]
    b.root(b.gate(G.OR, ev))
    g = b.build()
    exact = 1.0 - float(np.prod(1.0 - ps.mean()))
    n, n_outer = 1 << 15, 300
    u = uncertainty(g, Settings(num_trials=n, seed=5), ps, n_outer, deviate_seed=9, num_quantiles=10, num_bins=8)
    sigma_mean = u["std"][0] / math.sqrt(n_outer)
    assert abs(u["mean"][0] - exact) < 3 * sigma_mean, (u["mean"][0], exact, sigma_mean)
    assert u["ci_lo"][0] < u["mean"][0] < u["ci_hi"][0]
    # the spread is epistemic, not only binomial noise of the inner run
    assert u["std"][0] > 2 * math.sqrt(exact * (1 - exact) / n)
    q = u["quantiles"][0]
    srt = np.sort(u["p"][:, 0])
    assert np.all(np.diff(q) >= 0) and q[-1] == srt[-1] and q[4] == srt[n_outer // 2 - 1]   # level 0.5 = order statistic 150
    assert u["error_factor"][0] == pytest.approx(srt[math.ceil(0.95 * n_outer) - 1] / srt[n_outer // 2 - 1])
    assert int(u["bin_counts"][0].sum()) == n_outer and u["bin_edges"][0][0] == srt[0] and u["bin_edges"][0][-1] == srt[-1]
    hist, _ = np.histogram(u["p"][:, 0], bins=u["bin_edges"][0])
    assert np.array_equal(hist, u["bin_counts"][0].astype(np.int64))


def test_model_source_on_the_device_and_plan_restored():
    g = load_mef(text=UNC_XML, mission_time=1000.0)
    ps = g.meta["psource"]
    n = 8192 + 100
    u = uncertainty(g, Settings(num_trials=n, seed=2, tally_all_gates=True), ps, 4, deviate_seed=33)
    for t in range(4):
        p = np.clip(ps.sample(33, t), 2.0 ** -32, 1.0 - 2.0 ** -53)
        want = O.mc_scalar(with_p(g, p), 2, t * n, n, True)
        assert np.array_equal(u["p"][t] * n, want.astype(np.float64))
    # canopy_uncertainty_run leaves the plan at the mean probabilities
    import ctypes as C
    from canopy_b200 import _lib
    with Plan(g, tally_all_gates=True) as plan:
        before = plan.fingerprint()
        base, _ = plan.run(n, seed=2)
        o = plan._opts(2, 0, n, 0, 0, None, True, False)
        uo = _lib.CUncertaintyOpts()
        uo.n_trials, uo.seed = 3, 33
        stats = (_lib.CUncertaintyStats * plan.n_tallies)()
        _lib.check(_lib.lib().canopy_uncertainty_run(plan._h, ps._h, C.byref(o), C.byref(uo), stats, None, None, None, None))
        assert plan.fingerprint() == before
        again, _ = plan.run(n, seed=2)
        assert np.array_equal(base, again)
        assert stats[0].min <= stats[0].mean <= stats[0].max
