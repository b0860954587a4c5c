"""Scratch GPU probe: per-kernel times (profile pass, serialised) and the pipelined step for c4 / c2 with the loaded build."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
cfgs = sys.argv[1].split(",") if len(sys.argv) > 1 else ["c4", "c2"]
N = {"c4": 1 << 24, "c2": 1 << 26, "c3": 1 << 25, "c1": 1 << 26}
for cfg in cfgs:
    g = synth.config_graph(cfg) if cfg != "c1" else None
    n = N[cfg]
    with Plan(g, tally_all_gates=(cfg == "c3")) as plan:
        plan.run(n // 4, seed=1)
        _, p = plan.run(n, seed=1, profile=True)
        best = min(plan.run(n, seed=1)[1].total_ms for _ in range(2))
        t, r = plan.run(n, seed=1)
        print(json.dumps(dict(lib=os.path.basename(os.environ.get("CANOPY_LIB", "default")), cfg=cfg, sampler_ms=round(p.sampler_ms, 2),
                              gate_ms=round(p.gate_ms, 2), serial=round(p.sampler_ms + p.gate_ms, 2), pipelined=round(best, 2),
                              batches=r.n_batches, samples_per_s=round(n / best * 1e3), top=int(t[0]))), flush=True)
