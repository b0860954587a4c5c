#!/usr/bin/env python
"""bench.py -- headline benchmark of the Monte Carlo hot path (driver contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4] [--impl reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...       (one rank per GPU, NCCL)

A "step" is one pass of the hot path over the whole workload on every rank: rank r evaluates the
workload's S samples on the disjoint Philox counter range [r*S, (r+1)*S) (weak scaling), then ONE
all-reduce of the int64 tallies over NCCL.  `value` = N*S*K / max-over-ranks device time.
`e2e` goes through the public host API every step: lower the graph from HOST arrays (canopy_plan_create),
upload, run, all-reduce, and read the tallies back to the host.
The CPU arm (`cpu_baseline`, `--impl reference`) is the in-repo bit-packed multi-threaded port in oracle/
-- the reference itself has no implementation of this path to time.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mc_samples_per_sec"
UNIT = "samples/s"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        """Start polling (before the warm-up, so nvidia-smi is up when the timed region begins)."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def mark_begin(self):
        self.t0 = time.time()

    def stop(self):
        """Samples that arrived between mark_begin() and now (the timed region)."""
        t1 = time.time()
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for t, ln in self.lines:
            if t < getattr(self, "t0", 0.0) or t > t1 + 0.06:
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9 or not f[0].isdigit():
                continue
            try:
                sm.append(float(f[1])); smax = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def ncu_traffic(workload_name: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE gate_ring_kernel launch (592 slabs) from the committed
    `ncu --set full` capture of this workload under profiles/, in bytes; None if there is no capture for it."""
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    for tag in ("r01j", "r01i", "r01h"):      # newest evidence set first
        path = os.path.join(ROOT, "profiles", f"{tag}_ncu_full_gate_{workload_name}.csv")
        try:
            tot = 0.0
            for row in csv.reader(open(path)):
                if len(row) == 3 and row[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(row[2]) * scale[row[1]]
            if tot:
                return tot
        except Exception:
            continue
    return None


def workload(name: str):
    from canopy_b200 import synth
    if name == "c1":  # the hand-authored MEF example (the reference ships none), through the library's MEF reader
        from canopy_b200.engine import load_mef
        return load_mef(os.path.join(ROOT, "examples", "c1_small_ft.xml")), 1 << 20
    g = synth.config_graph(name)
    return g, synth.CONFIGS[name]["n_samples"]


def cpu_arm(graph, n_samples_hint: int, tally_all: bool, seconds_target: float = 12.0, spec: int = 2):
    """Bit-packed multi-threaded CPU port (oracle/) on a bounded sample of the same workload."""
    from oracle import oracle as O
    cores = len(os.sched_getaffinity(0))
    probe = 1 << 12
    t0 = time.perf_counter(); O.mc_packed(graph, 1, 0, probe, tally_all, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
    n = int(min(n_samples_hint, max(probe, probe * seconds_target / max(dt, 1e-6))))
    n = max(64, n // 64 * 64)
    t0 = time.perf_counter(); t = O.mc_packed(graph, 1, 0, n, tally_all, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
    return n / dt, cores, n, dt, t


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    g, S = workload(args.workload)
    tally_all = args.workload == "c3"
    from canopy_b200.engine import Plan
    stats = Plan(g, tally_all, sampler=args.sampler).stats  # host lowering only: sizes for the config block
    spec = stats["spec_version"]
    cores = len(os.sched_getaffinity(0))
    from oracle import oracle as O
    # bounded sample per step, sized from a probe so K+W steps end within a few minutes
    probe = 1 << 12
    t0 = time.perf_counter(); O.mc_packed(g, 1, 0, probe, tally_all, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
    budget = float(os.environ.get("CANOPY_BENCH_REF_BUDGET_S", "60")) / max(1, args.steps + args.warmup)  # seconds of CPU work per step
    n = int(max(probe, min(S, probe * budget / max(dt, 1e-6)))) // 64 * 64
    for w in range(args.warmup):
        O.mc_packed(g, 1, 0, n, tally_all, nthreads=cores, spec=spec)
    t0 = time.perf_counter()
    for k in range(args.steps):
        O.mc_packed(g, 1, (k + 1) * n, n, tally_all, nthreads=cores, spec=spec)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": config_block(args.workload, g, stats, S, n_step=n),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{n} samples/step of {args.workload} (bit-packed pthreads port in oracle/; the reference has no CPU path)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gate_evals_per_sec": v * stats["n_gates_live"]}
    print(json.dumps(line), flush=True)
    return 0


def config_block(name, g, stats, S, n_step=None):
    return {"workload": f"{name}: {'MEF example examples/c1_small_ft.xml' if name == 'c1' else 'synthetic fault tree'} B={g.n_basic} G={g.n_gates} (after DCE: B={stats['n_basic_live']} "
                        f"G={stats['n_gates_live']} E={stats['n_edges_live']} G'={stats['n_gates_stored']} layers={stats['n_layers']} "
                        f"max_fanin={stats['max_fanin']} mean_fanin={stats['mean_fanin']:.2f}), {S} samples/step/GPU"
                        + (f" (this arm: {n_step} samples/step)" if n_step else ""),
            "samples_per_step_per_gpu": n_step or S, "B_live": stats["n_basic_live"], "G_live": stats["n_gates_live"],
            "E_live": stats["n_edges_live"], "G_stored": stats["n_gates_stored"], "layers": stats["n_layers"],
            "spec_version": stats["spec_version"], "threshold_bits": 32,
            "l2": "no flush needed: each batch streams a value buffer of n_nodes*1 KB*batch_slabs (GBs) >> 126 MB L2"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="canopy_b200")
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4"])
    ap.add_argument("--sampler", type=int, default=0, choices=[0, 1, 2, 3],
                    help="sampling spec of the plan (0 = build default; include/canopy.h CANOPY_SAMPLER_*)")
    ap.add_argument("--no-study", action="store_true", help="skip the bounded c4 roofline study on rank 0")
    ap.add_argument("--samples", type=int, default=0,
                    help="samples per step per GPU instead of the workload's full count (side studies only; the line says so)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from canopy_b200.engine import Plan, microbench_int

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; canopy_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    g, S = workload(args.workload)
    if args.samples:
        S = args.samples
    tally_all = args.workload == "c3"
    plan = Plan(g, tally_all, sampler=args.sampler)
    stats = plan.stats
    nt = plan.n_tallies
    tallies = torch.zeros(max(nt, 1), dtype=torch.int64, device=dev)
    # A dedicated (non-default) torch stream: the library orders the whole run on the stream handle it is given,
    # so torch.cuda.Event on this stream brackets exactly the kernels of the timed steps.
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    def step(k, want_report=False):
        rep = plan.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=(k * world + rank) * S, device=local,
                              stream=stream, want_report=want_report)
        if world > 1:
            dist.all_reduce(tallies)  # the single collective: int64 tally sum over NVLink
        return rep

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local); clocks.start()
    rep0 = step(0, want_report=True)
    for w in range(1, args.warmup):
        step(w)
    launches_per_step = rep0.launches + (1 if world > 1 else 0)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    clocks.mark_begin()
    e0.record()
    for k in range(args.steps):
        step(args.warmup + k)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms = float(t_ms.item())
    value = world * S * args.steps / (ms * 1e-3)
    top_tally = int(tallies[0].item()) if nt else 0

    # ---- e2e: public host API, host buffers, H2D + D2H inside the timed region ------------------------
    e2e_steps = max(1, min(args.steps, 5))
    host_out = torch.zeros(max(nt, 1), dtype=torch.int64).pin_memory()

    def e2e_step(k):
        p2 = Plan(g, tally_all, sampler=args.sampler)  # host lowering from HOST arrays
        r = p2.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=(k * world + rank) * S, device=local,
                          stream=stream, want_report=True)   # uploads the CSR arrays (h2d) and runs
        if world > 1:
            dist.all_reduce(tallies)
        host_out.copy_(tallies, non_blocking=False)   # d2h of the step's result
        p2.close()
        return r
    e2e_step(0)
    barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        r_e2e = e2e_step(100 + k)
    barrier()
    dt = time.perf_counter() - t0
    t_e = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    e2e_value = world * S * e2e_steps / float(t_e.item())

    # ---- per-kernel roofline pass (rank 0 work replicated on all ranks to keep them in step) -----------
    prof = plan.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=0, device=local, stream=stream,
                           profile=True, want_report=True)
    peaks, peaks_src = load_peaks()
    alg_bytes_per_sample = (stats["n_edges_live"] + stats["n_gates_stored"]) / 8.0
    gate_launch_ms = prof.gate_ms / max(1, prof.gate_launches)
    gate_bytes_per_launch = alg_bytes_per_sample * S / max(1, prof.gate_launches)
    gate_gbs = gate_bytes_per_launch / (gate_launch_ms * 1e-3) / 1e9 if gate_launch_ms > 0 else 0.0
    sampler_bits = stats["n_basic_live"] * S
    sampler_rate = sampler_bits / (prof.sampler_ms * 1e-3) if prof.sampler_ms > 0 else 0.0

    line = None
    if rank == 0:
        # Integer roofline of the sampler.  Measured on sm_100 (canopy_microbench_int): LOP3, IMAD.WIDE.U32 and
        # any mix of the two issue at ONE shared rate of 64 lanes/clk/SM, so the bound is the integer instruction count.
        # Spec 2 with H = 7, per warp-row of 4096 samples (one 128-sample group per lane), from the ncu source page
        # profiles/r01g_ncu_source_sampler_c2_opcodes.csv: 135.1 IMAD.WIDE executed (8 Philox calls x 20 minus the
        # round-1..2 products the calls of a group share).  Algorithmic: each multiply pairs with one LOP3, + 2 LOP3 per
        # word and plane (56) + 4 = 330.2 warp-instr x 32 lanes / 4096 = 2.580 per Bernoulli bit.  Executed integer-class
        # (LOP3, IMAD*, ISETP, VIADD, SEL, SHF, ...): 453.5 per row = 3.543 per bit.  Spec 1: 12.3 per bit.
        int_peak, _ = microbench_int(0, local)
        spec = stats["spec_version"]
        kernel_name = {1: "sampler_kernel<32,1>", 2: "sampler_sliced_kernel<32,false>", 3: "sampler_sliced_kernel<32,true>"}[spec]
        if spec == 3:
            # Spec 3: events below 2^-5 take the sparse path.  Algorithmic integer work per 128-sample group (one lane): a
            # plane event 330.2 as above; a sparse event one Philox call (19 IMAD.WIDE -- the round-1 product of the event id
            # is warp-uniform -- + 20 LOP3) + a 64-bit compare (2) per table entry inspected (N + 1) + 4 per placed one
            # (extract 2, shift, or): 41 + 6 * 128 p.  Executed counts come from the ncu source page of this workload
            # (tools/sampler_opcounts.py -> profiles/sampler_opcounts.json), counted per warp instruction x 32 lanes.
            view = plan.view()
            thr = view["event_thr32"].astype(np.float64)
            sparse = thr < 2.0 ** 27
            per_group = np.where(sparse, 41.0 + 6.0 * 128.0 * thr / 2.0 ** 32, 330.2)
            alg_ops_per_bit = float(per_group.mean() / 128.0) if len(thr) else 0.0
            try:
                with open(os.path.join(ROOT, "profiles", "sampler_opcounts.json")) as f:
                    exec_ops_per_bit = float(json.load(f)["3"]["executed_int_ops_per_bit"]) if args.workload == "c2" else None
            except Exception:
                exec_ops_per_bit = None
            sampler_mix = {"sparse_events": int(sparse.sum()), "plane_events": int((~sparse).sum()),
                           "mean_ones_per_group_sparse": float((128.0 * thr[sparse] / 2.0 ** 32).mean()) if sparse.any() else 0.0}
        else:
            alg_ops_per_bit = 2.580 if spec == 2 else 581.0 * 2 / 128.0
            exec_ops_per_bit = 3.543 if spec == 2 else 12.3
            sampler_mix = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32", "data": "synthetic",
            "config": config_block(args.workload, g, stats, S),
            "gate_evals_per_sec": value * stats["n_gates_live"],
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(r_e2e.h2d_bytes + g.nbytes() * 0),
                    "d2h_bytes_per_step": int(nt * 8), "steps": e2e_steps,
                    "what": "canopy_plan_create from host arrays + upload + run + allreduce + tallies to host, wall clock"},
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": {"kernel": "gate_ring_kernel", "bound": "hbm", "achieved": gate_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": gate_gbs / peaks["hbm_gbs"], "traffic": ncu_traffic(args.workload),
                         "traffic_note": "DRAM read+write bytes of one full launch (2.42 M samples per launch: 16384 per SM) from the committed ncu --set full capture in profiles/; "
                                         "frac can read a little above 1: the peak is a 1:1 read:write copy, this kernel's traffic is 4:1 reads",
                         "alg_bytes_per_full_launch": alg_bytes_per_sample * 592 * 4096, "nominal_hbm_gbs": 7700.0, "frac_of_nominal": gate_gbs / 7700.0, "peak_source": peaks_src,
                         "alg_bytes_per_sample": alg_bytes_per_sample, "avg_launch_ms": gate_launch_ms,
                         "share_of_step": prof.gate_ms / max(prof.total_ms, 1e-9), "timed": "profile pass, kernels serialised"},
            "roofline_sampler": {"kernel": kernel_name,
                                 "bound": "int (LOP3 + IMAD.WIDE share one issue rate of 64 lanes/clk/SM)",
                                 "achieved": sampler_rate * alg_ops_per_bit / 1e9, "peak": int_peak / 1e9, "unit": "G int lane-instr/s",
                                 "frac": sampler_rate * alg_ops_per_bit / int_peak,
                                 "executed_frac": sampler_rate * exec_ops_per_bit / int_peak if exec_ops_per_bit else None,
                                 "event_mix": sampler_mix,
                                 "bernoulli_bits_per_s": sampler_rate, "alg_int_ops_per_bit": alg_ops_per_bit,
                                 "executed_int_ops_per_bit": exec_ops_per_bit,
                                 "avg_launch_ms": prof.sampler_ms / max(1, prof.sampler_launches),
                                 "share_of_step": prof.sampler_ms / max(prof.total_ms, 1e-9),
                                 "peak_source": "canopy_microbench_int(kind=0) measured in this run; instruction counts from the ncu source page in profiles/"},
            "top_event_tally_last_step": top_tally,
        }
    if world == 1 and rank == 0:
        cpu_v, cores, n_cpu, dt_cpu, _ = cpu_arm(g, S, tally_all, spec=stats["spec_version"])
        line["cpu_baseline"] = {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{n_cpu} samples of {args.workload} in {dt_cpu:.1f} s (bit-packed pthreads port in oracle/)"}
        if not args.no_study and args.workload != "c4":
            line["roofline_study_c4"] = c4_study(local, stream, peaks, sampler=args.sampler)
    if rank == 0:
        if args.samples:
            line["config"]["reduced"] = f"--samples {args.samples}: side study, NOT the workload's full sample count"
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def c4_study(local, stream, peaks, n=1 << 24, sampler=0):
    """Bounded pass over the 100k-event / 60k-gate DAG (BASELINE.json config 4): per-kernel numbers."""
    from canopy_b200 import synth
    from canopy_b200.engine import Plan
    g = synth.config_graph("c4")
    with Plan(g, sampler=sampler) as plan:
        st = plan.stats
        plan.run(n // 4, seed=7, device=local)
        _, full = plan.run(n, seed=7, device=local)
        _, prof = plan.run(n, seed=7, device=local, profile=True)
    b = (st["n_edges_live"] + st["n_gates_stored"]) / 8.0
    gbs = b * n / (prof.gate_ms * 1e-3) / 1e9
    return {"samples": n, "B_live": st["n_basic_live"], "G_live": st["n_gates_live"], "E_live": st["n_edges_live"],
            "layers": st["n_layers"], "max_fanin": st["max_fanin"],
            "samples_per_s_overlapped": n / (full.total_ms * 1e-3), "gate_evals_per_s": n / (full.total_ms * 1e-3) * st["n_gates_live"],
            "gate_kernel_GBs": gbs, "gate_kernel_frac_hbm": gbs / peaks["hbm_gbs"], "gate_ms": prof.gate_ms,
            "sampler_ms": prof.sampler_ms, "sampler_bits_per_s": st["n_basic_live"] * n / (prof.sampler_ms * 1e-3)}


if __name__ == "__main__":
    sys.exit(main())
