import json, os, sys
sys.path.insert(0, "/root/repo")
from canopy_b200 import synth
from canopy_b200.engine import Plan
g = synth.config_graph("c4")
n = 1 << 26
with Plan(g) as plan:
    for bs in (296, 370, 444, 148, 296):
        plan.run(n // 8, seed=1, batch_slabs=bs)
        best = min(plan.run(n, seed=1, batch_slabs=bs)[1].total_ms for _ in range(2))
        print(json.dumps(dict(batch_slabs=bs, ms=round(best, 2), samples_per_s=round(n / best * 1e3))), flush=True)
