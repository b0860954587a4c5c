import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
g = synth.config_graph("c2"); n = 1 << 24
plan = Plan(g); st = plan.stats
plan.run(n // 4, seed=1, profile=True)
t, rep = plan.run(n, seed=1, profile=True)
print(json.dumps(dict(pack=os.environ.get("CANOPY_SAMPLER_PACK"), sampler_ms=rep.sampler_ms, gate_ms=rep.gate_ms,
      bits_per_s=st["n_basic_live"] * n / (rep.sampler_ms * 1e-3), top=int(t[0]))))
