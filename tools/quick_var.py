"""Scratch GPU probe: run-to-run spread of the pipelined run over fresh plans (fresh streams)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
for cfg, n in (("c2", 1 << 25), ("c4", 1 << 23)):
    g = synth.config_graph(cfg)
    Plan(g).run(4096)
    out = []
    for it in range(10):
        with Plan(g) as p:
            p.run(n // 8)
            _, rep = p.run(n)
            out.append(round(rep.total_ms, 1))
    print(json.dumps(dict(cfg=cfg, conn=os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), ms=out)), flush=True)
