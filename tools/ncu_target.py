"""Short, deterministic workload for `ncu --set full` captures: two full batches of one config, no overlap."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
g = synth.config_graph(cfg)
with Plan(g) as plan:
    n = 592 * 4096 * 2
    t, rep = plan.run(n, seed=1, overlap=False, profile=True)
    print(cfg, "top", int(t[0]), "sampler_ms", rep.sampler_ms, "gate_ms", rep.gate_ms, "batches", rep.n_batches)
