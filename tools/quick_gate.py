"""Scratch GPU probe: gate-kernel timings for the bench configs (profile pass, kernels serialised)."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan

cfgs = sys.argv[1:] or ["c2", "c4"]
for cfg in cfgs:
    g = synth.config_graph(cfg)
    nsamp = (1 << 24) if cfg == "c2" else (1 << 22)
    with Plan(g, tally_all_gates=(cfg == "c3")) as plan:
        st = plan.stats
        plan.run(nsamp // 4, seed=1, profile=True, overlap=False)
        t, rep = plan.run(nsamp, seed=1, profile=True, overlap=False)
        b = (st["n_edges_live"] + st["n_gates_stored"]) / 8
        print(json.dumps(dict(cfg=cfg, kernel=os.environ.get("CANOPY_GATE_KERNEL", "ring"), ctas=os.environ.get("CANOPY_RING_CTAS"),
                              top=int(t[0]), gate_ms=rep.gate_ms, sampler_ms=rep.sampler_ms, batches=rep.n_batches,
                              gate_GBs=b * nsamp / (rep.gate_ms * 1e-3) / 1e9)), flush=True)
