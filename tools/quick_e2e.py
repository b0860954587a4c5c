"""Scratch GPU probe: where the end-to-end time of a one-shot quantification goes (host lowering, upload + buffer
allocation, run, teardown), the way bench.py's e2e leg calls it."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from canopy_b200 import synth
from canopy_b200.engine import Plan
dev = torch.device("cuda", 0)
ts = torch.cuda.Stream(device=dev); torch.cuda.set_stream(ts)
for cfg, n in (("c4", 1 << 25), ("c2", 1 << 26)):
    g = synth.config_graph(cfg)
    out = torch.zeros(8, dtype=torch.int64, device=dev)
    host = torch.zeros(8, dtype=torch.int64).pin_memory()
    p = Plan(g); p.run_device(out.data_ptr(), n, seed=1, device=0, stream=ts.cuda_stream, want_report=True); p.close()
    for it in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); p = Plan(g); t1 = time.perf_counter()
        rep = p.run_device(out.data_ptr(), n, seed=1, sample_begin=it * n, device=0, stream=ts.cuda_stream, want_report=True); t2 = time.perf_counter()
        host.copy_(out); t3 = time.perf_counter()
        p.close(); t4 = time.perf_counter()
        print(json.dumps(dict(cfg=cfg, lower_ms=(t1 - t0) * 1e3, run_wall_ms=(t2 - t1) * 1e3, run_dev_ms=rep.total_ms, d2h_ms=(t3 - t2) * 1e3,
                              close_ms=(t4 - t3) * 1e3, total_ms=(t4 - t0) * 1e3)), flush=True)
