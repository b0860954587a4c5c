"""Scratch GPU probe: integer-pipe microbenchmarks and per-kernel timings for the bench configs."""
import json, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canopy_b200 import synth
from canopy_b200.engine import Plan, microbench_int

out = {}
for kind, name in enumerate(["lop3", "imad_wide", "philox_round_mix", "popc"]):
    ops, ms = microbench_int(kind)
    out["peak_" + name] = ops
    print(f"{name}: {ops/1e12:.2f} T lane-instr/s ({ms:.3f} ms)", flush=True)

for cfg, nsamp, spec in (("c2", 1 << 24, 1), ("c2", 1 << 24, 2), ("c4", 1 << 22, 1), ("c4", 1 << 22, 2)):
    g = synth.config_graph(cfg)
    plan = Plan(g, sampler=spec)
    st = plan.stats
    for prof, ov in ((True, False), (False, False), (False, True)):
        plan.run(nsamp // 4, seed=1, profile=prof, overlap=ov)
        t, rep = plan.run(nsamp, seed=1, profile=prof, overlap=ov)
        sps = nsamp / (rep.total_ms * 1e-3)
        line = dict(cfg=cfg, spec=spec, profile=prof, overlap=ov, n=nsamp, total_ms=rep.total_ms, sampler_ms=rep.sampler_ms, gate_ms=rep.gate_ms,
                    samples_per_s=sps, batches=rep.n_batches, batch_slabs=rep.batch_slabs, top=int(t[0]))
        if prof and rep.gate_ms > 0:
            bytes_per_sample = (st["n_edges_live"] + st["n_gates_stored"]) / 8
            line["gate_GBs"] = bytes_per_sample * nsamp / (rep.gate_ms * 1e-3) / 1e9
            line["sampler_bits_per_s"] = st["n_basic_live"] * nsamp / (rep.sampler_ms * 1e-3)
        print(json.dumps(line), flush=True)
    plan.close()
