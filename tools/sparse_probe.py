"""Scratch GPU probe for spec 3 tuning builds (CANOPY_LIB=...): sampler / gate ms (profile pass) and pipelined ms on c2 and c4."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan

spec = int(sys.argv[1]) if len(sys.argv) > 1 else 3
only = sys.argv[2] if len(sys.argv) > 2 else None
for cfg, n in (("c2", 1 << 25), ("c4", 1 << 22)):
    if only and cfg != only:
        continue
    g = synth.config_graph(cfg)
    with Plan(g, sampler=spec) as plan:
        plan.run(n // 4, seed=1)
        _, prof = plan.run(n, seed=1, profile=True)
        best = min(plan.run(n, seed=1)[1].total_ms for _ in range(3))
        print(json.dumps(dict(lib=os.path.basename(os.environ.get("CANOPY_LIB", "default")), cfg=cfg, spec=spec, n=n,
                              sampler_ms=round(prof.sampler_ms, 3), gate_ms=round(prof.gate_ms, 3), pipelined_ms=round(best, 3),
                              samples_per_s=n / (best * 1e-3))), flush=True)
