"""Scratch GPU probe: sampler time per probability class (the c2 graph with every event set to one probability), to see
which path is latency-bound at reduced occupancy (run with the default build and with a low-occupancy variant)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from canopy_b200 import synth
from canopy_b200.engine import Plan
n = 1 << 24
for label, p in (("dense p=0.1", 0.1), ("k=7 p=0.02", 0.02), ("k=9 p=0.005", 0.005), ("k=11 p=0.0012", 0.0012), ("k=13 p=0.0002", 0.0002)):
    g = synth.config_graph("c2")
    g.p[:] = p
    with Plan(g, tally_all_gates=True) as plan:
        plan.run(n // 4, seed=1)
        _, r = plan.run(n, seed=1, profile=True)
        print(json.dumps(dict(lib=os.path.basename(os.environ.get("CANOPY_LIB", "default")), cls=label, events=int(plan.stats["n_basic_live"]),
                              sampler_ms=round(r.sampler_ms, 3), ns_per_row=round(r.sampler_ms * 1e6 / (plan.stats["n_basic_live"] * n / 8192), 2))), flush=True)
