"""Scratch GPU probe: cost of one epistemic trial of canopy_uncertainty_run (draw all p on the host, re-quantise, rebuild the
count tables on the device, inner run, tallies to host) against the inner Monte Carlo run alone."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from canopy_b200 import synth
from canopy_b200.engine import DEV_LOGNORMAL_EF, Plan, PSource, Settings, uncertainty
for cfg, n in (("c2", 1 << 22), ("c4", 1 << 22)):
    g = synth.config_graph(cfg)
    ps = PSource.from_deviates([(DEV_LOGNORMAL_EF, float(p), 3.0, 0.95) for p in g.p])
    t0 = time.perf_counter(); [ps.sample(1, t) for t in range(8)]; draw_ms = (time.perf_counter() - t0) / 8 * 1e3
    with Plan(g) as plan:
        plan.run(n, seed=1)
        t0 = time.perf_counter(); plan.run(n, seed=1, sample_begin=n); inner_ms = (time.perf_counter() - t0) * 1e3
    trials = 16
    uncertainty(g, Settings(num_trials=n, seed=1), ps, 2, deviate_seed=5)
    t0 = time.perf_counter(); u = uncertainty(g, Settings(num_trials=n, seed=1), ps, trials, deviate_seed=5); dt = time.perf_counter() - t0
    print(json.dumps(dict(cfg=cfg, events=g.n_basic, inner_samples=n, inner_run_ms=round(inner_ms, 2), draw_ms=round(draw_ms, 2),
                          ms_per_trial=round(dt / trials * 1e3, 2), mean=float(u["mean"][0]), std=float(u["std"][0]),
                          error_factor=float(u["error_factor"][0]))), flush=True)
