"""Short workload for an `ncu --set full` capture of the fused engine: config c1 (the MEF example), 2^20 samples."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200.engine import Plan, load_mef
g = load_mef(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "c1_small_ft.xml"))
with Plan(g, tally_all_gates=True) as plan:
    plan.run(1 << 20, seed=1)
    t, rep = plan.run(1 << 20, seed=1)
    print("c1 top", int(t[0]), "ms", rep.total_ms, "launches", rep.gate_launches, rep.sampler_launches)
