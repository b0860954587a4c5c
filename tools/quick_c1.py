"""Scratch GPU probe: where a small run (config c1, 2^20 samples) spends its time."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200.engine import Plan, load_mef
g = load_mef(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "c1_small_ft.xml"))
with Plan(g) as p:
    print(p.stats)
    p.run(1 << 20)
    for n in (1 << 20, 1 << 22, 1 << 24):
        t0 = time.perf_counter(); _, rep = p.run(n); t1 = time.perf_counter()
        _, prof = p.run(n, profile=True)
        print(json.dumps(dict(n=n, wall_ms=(t1 - t0) * 1e3, dev_ms=rep.total_ms, batches=rep.n_batches, batch_slabs=rep.batch_slabs,
                              sampler_ms=prof.sampler_ms, gate_ms=prof.gate_ms, prof_total=prof.total_ms)))
