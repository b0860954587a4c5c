"""Scratch GPU probe: pipelined ms on c2 (2^26 samples) and c4 (2^23) against batch size; CANOPY_RING_CTAS etc. come from the environment."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan
for cfg, n, sizes in (("c2", 1 << 26, (0, 148, 592, 1184)), ("c4", 1 << 23, (0, 148))):
    g = synth.config_graph(cfg)
    with Plan(g) as plan:
        plan.run(n // 8, seed=1)
        for bs in sizes:
            plan.run(n // 8, seed=1, batch_slabs=bs)
            best = min(plan.run(n, seed=1, batch_slabs=bs)[1].total_ms for _ in range(2))
            print(json.dumps(dict(env={k: v for k, v in os.environ.items() if k.startswith("CANOPY_")}, cfg=cfg, batch_slabs=bs,
                                  ms=round(best, 3), samples_per_s=n / (best * 1e-3))), flush=True)
