"""Scratch GPU probe: where the end-to-end time of a one-shot quantification goes (host lowering, upload + buffer
allocation, run, teardown)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from canopy_b200 import synth
from canopy_b200.engine import Plan
for cfg, n in (("c2", 1 << 26), ("c4", 1 << 23)):
    g = synth.config_graph(cfg)
    Plan(g).run(4096)  # context + module load
    for it in range(3):
        t0 = time.perf_counter(); p = Plan(g); t1 = time.perf_counter()
        p.run(4096); t2 = time.perf_counter()          # device state + plan upload (tiny run)
        _, rep = p.run(n); t3 = time.perf_counter()     # value-buffer allocation + the run
        _, rep2 = p.run(n); t4 = time.perf_counter()    # steady state
        p.close(); t5 = time.perf_counter()
        print(json.dumps(dict(cfg=cfg, lower_ms=(t1 - t0) * 1e3, state_ms=(t2 - t1) * 1e3, first_run_ms=(t3 - t2) * 1e3,
                              first_dev_ms=rep.total_ms, steady_ms=(t4 - t3) * 1e3, steady_dev_ms=rep2.total_ms, close_ms=(t5 - t4) * 1e3)), flush=True)
