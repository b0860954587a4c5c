"""Scratch GPU probe: whole-run device time with and without sampler/gate overlap (env knobs set by the caller)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
for cfg, n in (("c2", 1 << 25), ("c4", 1 << 23)):
    g = synth.config_graph(cfg)
    with Plan(g) as plan:
        res = {}
        for ov in (False, True):
            plan.run(n // 4, seed=1, overlap=ov)
            t, rep = plan.run(n, seed=1, overlap=ov)
            res["overlap" if ov else "serial"] = round(rep.total_ms, 2)
        print(json.dumps(dict(cfg=cfg, smem=os.environ.get("CANOPY_SAMPLER_SMEM"), ring_ctas=os.environ.get("CANOPY_RING_CTAS"),
                              top=int(t[0]), **res, gain=round(res["serial"] / res["overlap"], 3))), flush=True)
