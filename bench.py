#!/usr/bin/env python
"""bench.py -- headline benchmark of the Monte Carlo hot path (driver contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c3|c1] [--impl reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...       (one rank per GPU, NCCL)

Default workload: BASELINE.json config 4 / 5, the 100k-event / 60k-gate deep DAG with fan-in up to 256 (the
configuration north_star states its target on).  The config asks for 2^30 samples on one GPU (2^36 over eight): one
step here is 2^27 samples per GPU, because the driver times 25 steps per arm, four GPU counts and the CPU arm in one
lease and 2^30 per step would take minutes per step count; the graph, and therefore bytes and instructions per
sample, are the full configuration's (a step is 55 full batches of the pipeline, fill and drain are below 2 %).
tools/run_c5.py runs the 2^36-sample configuration in one go.

A "step" is one pass of the hot path over the workload's S samples on every rank: rank r evaluates the disjoint
Philox counter range [(k*N + r)*S, +S) (weak scaling) into row k of a [steps, n_tallies] int64 tensor; after the K timed
steps ONE all-reduce over NCCL sums those rows across the ranks, inside the timed region (north_star: "a single NCCL
allreduce of int64 tallies"; until round 2 every step had its own all-reduce, which re-synchronised the ranks K times).
`value` = N*S*K / max-over-ranks device time; `rank_ms` lists each rank's own device time per step.  Before the timed region `tally_check` runs a FIXED range cut over the
N ranks by canopy_shard_range with every gate tallied, all-reduces it, and compares it on rank 0 with the same range on
one GPU and (a prefix) with the CPU oracle; a mismatch exits non-zero.
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


def build_tag() -> str:
    """Hash of the kernel sources: profiles/ captures carry the tag of the build they were taken from, and a number read
    from a capture of ANOTHER build is reported as null instead of silently going stale."""
    import hashlib
    h = hashlib.sha1()
    for f in ("engine.cu", "kernels.cuh", "sampler.cuh", "walk.cuh", "small.cuh", "sparse_table.h", "plan.h", "lowering.cpp"):
        with open(os.path.join(ROOT, "canopy_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:12]


def captured(workload_name: str):
    """profiles/captures.json: per workload the DRAM bytes of one gate-walk launch (ncu --set full,
    dram__bytes_read.sum + dram__bytes_write.sum) and the executed integer instructions per Bernoulli bit of the sampler
    (ncu source page), each with the build tag it was measured on (tools/gpu_evidence.sh writes it)."""
    try:
        with open(os.path.join(ROOT, "profiles", "captures.json")) as f:
            c = json.load(f).get(workload_name)
    except Exception:
        return None, None, "no capture in profiles/captures.json"
    if not c:
        return None, None, "no capture for this workload in profiles/captures.json"
    if c.get("build_tag") != build_tag():
        return None, None, f"capture is of build {c.get('build_tag')}, this build is {build_tag()}: not quoted"
    return c.get("gate_dram_bytes_per_launch"), c.get("sampler_executed_int_ops_per_bit_x_events"), f"ncu capture {c.get('capture')} of build {c['build_tag']}"


# samples per step per GPU where the configuration's full count makes a step too long for 25 timed steps per arm
STEP_SAMPLES = {"c4": 1 << 27}


def _c1_graph_stdlib(path):
    """The c1 example (AND/OR gates over basic events with float probabilities) read with xml.etree: the CPU arm must not
    load libcanopy.so, not even for parsing.  Same node order as the library's reader (tests/test_bench_contract.py)."""
    import xml.etree.ElementTree as ET
    from canopy_b200.graph import AND, OR, GraphBuilder
    root = ET.parse(path).getroot()
    gates = [(d.get("name"), d[0]) for d in root.iter("define-gate")]
    used = {a.get("name") for _, f in gates for a in f if a.tag == "gate"}
    basics = [(d.get("name"), float(d.find("float").get("value"))) for d in root.iter("define-basic-event")]
    b = GraphBuilder()
    ev = {n: b.basic_event(p, n) for n, p in basics}
    gh = {n: ("g", i) for i, (n, _) in enumerate(gates)}
    for n, f in gates:
        b.gate({"and": AND, "or": OR}[f.tag], [gh[a.get("name")] if a.tag == "gate" else ev[a.get("name")] for a in f], name=n)
    for n, _ in gates:
        if n not in used:
            b.root(gh[n])
    return b.build()


def workload(name: str, native: bool = True):
    """(graph, samples per step per GPU, samples of the full configuration)"""
    from canopy_b200 import synth
    if name == "c1":  # the hand-authored MEF example (the reference ships none)
        path = os.path.join(ROOT, "examples", "c1_small_ft.xml")
        if native:   # through the library's MEF reader
            from canopy_b200.engine import load_mef
            return load_mef(path), 1 << 20, 1 << 20
        return _c1_graph_stdlib(path), 1 << 20, 1 << 20
    g = synth.config_graph(name)
    full = synth.CONFIGS[name]["n_samples"]
    return g, STEP_SAMPLES.get(name, full), full


WHAT = {"c1": "MEF example examples/c1_small_ft.xml", "c2": "synthetic fault tree incl. ATLEAST and NOT",
        "c3": "synthetic event tree, 200 sequences over one shared DAG, per-sequence and per-gate tallies",
        "c4": "deep synthetic DAG, fan-in up to 256 (BASELINE configs 4 and 5)"}


def config_block(name, g, S, S_full, stats=None, n_step=None):
    """`workload` is built from the DESCRIPTOR only, so both arms print the same string; the GPU arm adds the sizes after
    dead-code elimination from the plan."""
    import numpy as np
    fan = np.diff(g.child_off.astype(np.int64))
    c = {"workload": f"{name}: {WHAT[name]}; B={g.n_basic} G={g.n_gates} E={g.n_edges} max_fanin={int(fan.max()) if len(fan) else 0}; "
                     f"{S} samples/step/GPU" + (f" (the configuration's {S_full} samples = {S_full // S} such steps)" if S_full != S else ""),
         "samples_per_step_per_gpu": n_step or S, "threshold_bits": 32,
         "l2": "no flush needed: each batch streams a value buffer of n_nodes*1 KB*batch_slabs (GBs) >> 126 MB L2"}
    if n_step and n_step != S:
        c["this_arm"] = f"{n_step} samples/step (bounded CPU sample of the same workload)"
    if stats:
        c.update({"B_live": stats["n_basic_live"], "G_live": stats["n_gates_live"], "E_live": stats["n_edges_live"],
                  "G_stored": stats["n_gates_stored"], "layers": stats["n_layers"], "max_fanin_live": stats["max_fanin"],
                  "mean_fanin_live": stats["mean_fanin"], "spec_version": stats["spec_version"]})
    return c


def cpu_arm(graph, n_samples_hint: int, tally_all: bool, seconds_target: float = 12.0, spec: int = 0, bits: int = 32):
    """Bit-packed multi-threaded CPU port (oracle/) on a bounded sample of the same workload."""
    from oracle import oracle as O
    spec = spec or O.SPEC
    cores = len(os.sched_getaffinity(0))
    # size the sample from probes that are long enough to mean something (thread start-up dominates a 4096-sample call)
    probe, dt = 1 << 12, 0.0
    while True:
        t0 = time.perf_counter(); O.mc_packed(graph, 1, 0, probe, tally_all, bits=bits, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
        if dt >= 0.5 or probe >= n_samples_hint:
            break
        probe *= 4
    n = int(min(n_samples_hint, max(probe, probe * seconds_target / max(dt, 1e-6))))
    n = max(64, n // 64 * 64)
    t0 = time.perf_counter(); t = O.mc_packed(graph, 1, 0, n, tally_all, bits=bits, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
    return n / dt, cores, n, dt, t


def run_reference(args):
    """The CPU arm on the driver's clock.  Nothing of canopy_b200's native code is loaded here: the graph generator is
    pure Python, the timed loop is oracle/ only (the reference has no implementation of this path to time)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    g, S, S_full = workload(args.workload, native=False)
    tally_all = args.workload == "c3"
    cores = len(os.sched_getaffinity(0))
    from oracle import oracle as O
    spec = args.sampler or O.SPEC
    bits = args.threshold_bits
    # bounded sample per step, sized from a probe so K+W steps end within a few minutes
    probe, dt = 1 << 12, 0.0
    while True:
        t0 = time.perf_counter(); O.mc_packed(g, 1, 0, probe, tally_all, bits=bits, nthreads=cores, spec=spec); dt = time.perf_counter() - t0
        if dt >= 0.5 or probe >= S:
            break
        probe *= 4
    budget = float(os.environ.get("CANOPY_BENCH_REF_BUDGET_S", "60")) / max(1, args.steps + args.warmup)  # seconds of CPU work per step
    n = int(max(probe, min(S, probe * budget / max(dt, 1e-6)))) // 64 * 64
    for w in range(args.warmup):
        O.mc_packed(g, 1, 0, n, tally_all, bits=bits, nthreads=cores, spec=spec)
    t0 = time.perf_counter()
    for k in range(args.steps):
        O.mc_packed(g, 1, (k + 1) * n, n, tally_all, bits=bits, nthreads=cores, spec=spec)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": f"u{bits}", "data": "synthetic",
            "config": dict(config_block(args.workload, g, S, S_full, n_step=n), threshold_bits=bits),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{n} samples/step of {args.workload} (bit-packed pthreads port in oracle/, spec {spec}; the reference has no CPU path)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gate_evals_per_sec_descriptor_gates": v * g.n_gates}
    print(json.dumps(line), flush=True)
    return 0


CHECK_SEED = 0x7A11


def tally_check(g, world, rank, local, dev, stream, sampler, n_fixed=1 << 24, n_prefix=None, bits=32, host_group=None):
    """In-run parity evidence on the driver's clock, every gate tallied:
       (i)  the fixed range [0, n_fixed) cut over the N ranks by canopy_shard_range + ONE all-reduce == the same range on
            one GPU (at N = 1: three ragged pieces summed == the whole);
       (ii) its first n_prefix samples on the GPU == oracle.mc_packed on the host; by default one full batch of the
            default pipeline (two slabs per SM) plus a slab and a ragged tail, so that the run the oracle checks crosses a
            swap of the two value buffers (c4: 2.4e6 samples, about 10 s of host time)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from canopy_b200.engine import Plan, shard_range
    if n_prefix is None:
        n_prefix = (2 * torch.cuda.get_device_properties(local).multi_processor_count + 1) * 8192 + 77
    res = {"range": [0, n_fixed], "ranks": world, "prefix": n_prefix, "tallies": 0}
    with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=sampler) as plan:
        nt = plan.n_tallies
        res["tallies"] = nt
        buf = torch.zeros(max(nt, 1), dtype=torch.int64, device=dev)
        b, n = shard_range(0, n_fixed, rank, world)
        plan.run_device(buf.data_ptr(), n, seed=CHECK_SEED, sample_begin=b, device=local, stream=stream)
        if world > 1:
            dist.all_reduce(buf)      # the collective under test: nt int64 tallies (c4: 460 KB)
        torch.cuda.synchronize()
        flags = torch.ones(3, dtype=torch.int64, device=dev)
        if rank == 0:
            nway = buf.cpu().numpy().astype(np.uint64)[:nt]
            one, _ = plan.run(n_fixed, seed=CHECK_SEED, sample_begin=0, device=local)
            flags[0] = int(np.array_equal(one, nway))
            cuts = [0, 4096 * 301 + 77, n_fixed // 2 + 12345, n_fixed]
            pieces = sum(plan.run(cuts[i + 1] - cuts[i], seed=CHECK_SEED, sample_begin=cuts[i], device=local)[0] for i in range(3))
            flags[1] = int(np.array_equal(one, pieces))
            from oracle import oracle as O
            pre, _ = plan.run(n_prefix, seed=CHECK_SEED, sample_begin=0, device=local)
            want = O.mc_packed(g, CHECK_SEED, 0, n_prefix, True, bits=bits, nthreads=len(os.sched_getaffinity(0)), spec=plan.spec)
            flags[2] = int(np.array_equal(pre, want))
            res["top_event_tally"] = int(one[0]) if nt else 0
            res["nonzero_tallies"] = int((one > 0).sum())
        if world > 1:
            dist.broadcast(flags, 0)
        f = [bool(x) for x in flags.cpu().tolist()]
    res.update({"n_way_equals_1_way": f[0], "pieces_equal_whole": f[1], "prefix_equals_oracle": f[2],
                "allreduce_bytes": nt * 8 if world > 1 else 0})
    if world > 1 and host_group is not None:
        # (iii) the library's own collective: canopy_mc_run_multi (csrc/multi.cu: ncclCommInitAll + one
        # ncclAllReduce(ncclUint64, ncclSum)) driven from rank 0's process over all `world` GPUs, against the same range on
        # one GPU.  Every rank has closed its check plan; the others wait on a HOST (gloo) barrier, so that no kernel of
        # theirs spins on the GPUs rank 0 is about to use.
        torch.cuda.synchronize()
        multi = torch.ones(1, dtype=torch.int64, device=dev)
        if rank == 0:
            n_multi = 1 << 22
            t0 = time.perf_counter()
            with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=sampler) as plan:
                many, _ = plan.run_multi(list(range(world)), n_multi, seed=CHECK_SEED, sample_begin=12345, batch_slabs=32)
                one, _ = plan.run(n_multi, seed=CHECK_SEED, sample_begin=12345, device=local, batch_slabs=32)
            multi[0] = int(np.array_equal(many, one))
            res["run_multi"] = {"devices": world, "samples": n_multi, "seconds": round(time.perf_counter() - t0, 2),
                                "what": "canopy_mc_run_multi from rank 0's process (own ncclCommInitAll + ncclAllReduce) == one GPU"}
        dist.barrier(group=host_group)
        dist.broadcast(multi, 0)
        res["run_multi_equals_1_way"] = bool(multi.item())
    return res


def sampler_algorithmic_ops(plan, spec):
    """Algorithmic integer lane-instructions per Bernoulli bit of the sampler for this plan's event mix."""
    import numpy as np
    if spec == 1:
        return 581.0 * 2 / 128.0, None
    if spec == 2:
        return 2.580, None
    view = plan.view()
    thr = view["event_thr64"].astype(np.float64) / 2.0 ** 64
    k = view["event_k"].astype(np.int64)      # 0 = dense (planes), 7..13 = sparse with 2^k-sample groups
    sparse = k > 0
    grp = np.where(sparse, 2.0 ** k, 128.0)
    # per group (one lane): a dense event 330.2 per 128 samples; a sparse event one Philox call
    # (19 IMAD.WIDE + 20 LOP3) + a 64-bit compare (2) per table entry inspected (N + 1) + 4 per placed one: 41 + 6 * 2^k p
    per_group = np.where(sparse, 41.0 + 6.0 * grp * thr, 330.2)
    mix = {"sparse_events": int(sparse.sum()), "plane_events": int((~sparse).sum()),
           "events_per_class_k7_to_k13": [int((k == c).sum()) for c in range(7, 14)],
           "mean_ones_per_group_sparse": float((grp[sparse] * thr[sparse]).mean()) if sparse.any() else 0.0}
    return (float((per_group / grp).mean()) if len(thr) else 0.0), mix


def side_line(name, local, stream, sampler, steps=2, bits=32, samples=None, warmups=2):
    """A short device-timed line of another BASELINE config on rank 0 (extra key, not the headline)."""
    import torch
    from canopy_b200.engine import Plan
    g, S, S_full = workload(name)
    S = samples or S
    tally_all = name == "c3"
    with Plan(g, tally_all, threshold_bits=bits, sampler=sampler) as plan:
        nt = plan.n_tallies
        out = torch.zeros(max(nt, 1), dtype=torch.int64, device=torch.device("cuda", local))
        for w in range(warmups):
            plan.run_device(out.data_ptr(), S, seed=0xB200, sample_begin=w * S, device=local, stream=stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            plan.run_device(out.data_ptr(), S, seed=0xB200, sample_begin=(2 + k) * S, device=local, stream=stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        st = plan.stats
    return {"workload": config_block(name, g, S, S_full)["workload"], "threshold_bits": bits, "samples_per_s": S / (ms * 1e-3), "ms_per_step": ms, "steps": steps,
            "gate_evals_per_s": S / (ms * 1e-3) * st["n_gates_live"], "tallies": nt, "G_live": st["n_gates_live"],
            "spec_version": st["spec_version"], "samples_per_step": S, "warmup_steps": warmups,
            "alg_bytes_per_sample": (st["n_edges_live"] + st["n_gates_stored"] + st["n_basic_live"]) / 8.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="canopy_b200")
    ap.add_argument("--workload", default="c4", choices=["c1", "c2", "c3", "c4"])
    ap.add_argument("--sampler", type=int, default=0,
                    help="sampling spec of the plan (0 = build default; include/canopy.h CANOPY_SAMPLER_*)")
    ap.add_argument("--threshold-bits", type=int, default=32, choices=[32, 64],
                    help="width of the variate / threshold comparison (canopy_plan_opts.threshold_bits)")
    ap.add_argument("--no-study", action="store_true", help="skip the short c2 / c3 side lines on rank 0")
    ap.add_argument("--no-check", action="store_true", help="skip tally_check (kernel-tuning runs only; the line says so)")
    ap.add_argument("--samples", type=int, default=0,
                    help="samples per step per GPU instead of the workload's count (side studies only; the line says so)")
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
    host_group = dist.new_group(backend="gloo") if world > 1 else None   # host-side barrier only (tally_check iii)
    dev = torch.device("cuda", local)

    g, S, S_full = workload(args.workload)
    if args.samples:
        S = args.samples
    tally_all = args.workload == "c3"
    # A dedicated (non-default) torch stream: the library orders the whole run on the stream handle it is given,
    # so torch.cuda.Event on this stream brackets exactly the kernels of the timed steps.
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    # ---- parity evidence before anything is timed ------------------------------------------------------
    check = None
    if not args.no_check:
        check = tally_check(g, world, rank, local, dev, stream, args.sampler, bits=args.threshold_bits, host_group=host_group)
        if not (check["n_way_equals_1_way"] and check["pieces_equal_whole"] and check["prefix_equals_oracle"]
                and check.get("run_multi_equals_1_way", True)):
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "tally_check failed", "tally_check": check}), flush=True)
            if world > 1:
                dist.destroy_process_group()
            return 3

    plan = Plan(g, tally_all, threshold_bits=args.threshold_bits, sampler=args.sampler)
    stats = plan.stats
    spec = stats["spec_version"]
    nt = plan.n_tallies
    tallies = torch.zeros(max(nt, 1), dtype=torch.int64, device=dev)

    # One row of tallies per step; ONE collective over the rows of a run after its last step (north_star: "a single NCCL
    # allreduce of int64 tallies").  Round 2 reduced after every step, which re-synchronised the ranks K times per run.
    K, W = args.steps, args.warmup
    rows = torch.zeros((W + K, max(nt, 1)), dtype=torch.int64, device=dev)

    def step(k, want_report=False):
        return plan.run_device(rows[k].data_ptr(), S, seed=0xB200, sample_begin=(k * world + rank) * S, device=local,
                               stream=stream, want_report=want_report)

    def reduce_rows(a, b):
        if world > 1:
            dist.all_reduce(rows[a:b])  # the single collective: int64 tally sum over NVLink

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local); clocks.start()
    rep0 = step(0, want_report=True)
    for w in range(1, W):
        step(w)
    reduce_rows(0, W)
    barrier()
    e0, e_own, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    barrier()
    clocks.mark_begin()
    e0.record()
    for k in range(K):
        step(W + k)
    e_own.record()     # this rank's own K steps end here; what follows is the collective (and the wait for the slowest rank)
    reduce_rows(W, W + K)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    own_ms = e0.elapsed_time(e_own)
    clk = clocks.stop()
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    rank_ms = None
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        mine = torch.tensor([own_ms / K], dtype=torch.float64, device=dev)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        per_rank = [float(x.item()) for x in every]
        rank_ms = {"per_rank_ms_per_step": per_rank, "min": min(per_rank), "median": statistics.median(per_rank), "max": max(per_rank),
                   "allreduce_and_wait_ms_rank0": ms - own_ms,
                   "what": "device ms per step of each rank's own K steps (CUDA events before the collective); the job's step time is the max"}
    ms = float(t_ms.item())
    value = world * S * K / (ms * 1e-3)
    top_tally = int(rows[W + K - 1][0].item()) if nt else 0
    launches_total = int(rep0.launches * K + (1 if world > 1 else 0))

    # ---- per-kernel roofline pass (every rank, to keep them in step) --------------------------------------
    prof = plan.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=0, device=local, stream=stream,
                           profile=True, want_report=True)
    # the same pass with every walk launched in the shape the timed steps give it (one CTA per SM), still alone
    prof_shape = plan.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=0, device=local, stream=stream,
                                 profile=True, want_report=True, profile_walk_shape=1)
    alg_ops_per_bit, sampler_mix = sampler_algorithmic_ops(plan, spec)
    plan.close()   # its value buffers (c4: 2 x 48 GB) go to the library's pool: the e2e plans below reuse them

    # ---- e2e: public host API, host buffers, H2D + D2H inside the timed region ------------------------
    e2e_steps = max(1, min(args.steps, 5))
    host_out = torch.zeros(max(nt, 1), dtype=torch.int64).pin_memory()

    e2e_parts = {"plan_create_ms": 0.0, "run_ms": 0.0, "run_device_ms": 0.0, "allreduce_d2h_ms": 0.0, "plan_destroy_ms": 0.0}

    def e2e_step(k):
        t0 = time.perf_counter()
        p2 = Plan(g, tally_all, threshold_bits=args.threshold_bits, sampler=args.sampler)  # canopy_plan_create from HOST arrays
        t1 = time.perf_counter()
        r = p2.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=(k * world + rank) * S, device=local,
                          stream=stream, want_report=True)   # uploads the lowered arrays (h2d) and runs
        t2 = time.perf_counter()
        if world > 1:
            dist.all_reduce(tallies)
        host_out.copy_(tallies, non_blocking=False)   # d2h of the step's result
        t3 = time.perf_counter()
        p2.close()
        t4 = time.perf_counter()
        for key, v in zip(e2e_parts, (t1 - t0, t2 - t1, r.total_ms * 1e-3, t3 - t2, t4 - t3)):
            e2e_parts[key] += v * 1e3
        return r
    e2e_step(0)
    barrier()
    for key in e2e_parts:
        e2e_parts[key] = 0.0
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        r_e2e = e2e_step(100 + k)
    barrier()
    dt = time.perf_counter() - t0
    t_e = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    e2e_value = world * S * e2e_steps / float(t_e.item())

    # ---- cold first call (N = 1): what a one-shot user of quantify() pays -- no plan cache, no pooled buffers ------
    e2e_cold = None
    if world == 1:
        from canopy_b200.engine import plan_cache_clear, trim_device_cache
        torch.cuda.synchronize()
        trim_device_cache(-1); plan_cache_clear()
        t0 = time.perf_counter()
        p3 = Plan(g, tally_all, threshold_bits=args.threshold_bits, sampler=args.sampler, plan_cache=False)
        t1 = time.perf_counter()
        p3.run_device(tallies.data_ptr(), S, seed=0xB200, sample_begin=0, device=local, stream=stream, want_report=True)
        host_out.copy_(tallies, non_blocking=False)
        t2 = time.perf_counter()
        p3.close()
        t3 = time.perf_counter()
        e2e_cold = {"value": S / (t3 - t0), "unit": UNIT, "steps": 1, "plan_create_ms": (t1 - t0) * 1e3, "run_ms": (t2 - t1) * 1e3,
                    "plan_destroy_ms": (t3 - t2) * 1e3,
                    "what": "one step after canopy_trim_device_cache + canopy_plan_cache_clear with no_plan_cache = 1: lowering from "
                            "scratch, cudaMalloc of the plan arena and both value buffers, run, tallies to host, destroy"}

    peaks, peaks_src = load_peaks()
    gate_bytes_per_sample = (stats["n_edges_live"] + stats["n_gates_stored"]) / 8.0
    sampler_bytes_per_sample = stats["n_basic_live"] / 8.0
    gate_launch_ms = prof.gate_ms / max(1, prof.gate_launches)
    gate_bytes_per_launch = gate_bytes_per_sample * S / max(1, prof.gate_launches)
    gate_gbs = gate_bytes_per_launch / (gate_launch_ms * 1e-3) / 1e9 if gate_launch_ms > 0 else 0.0
    sampler_bits = stats["n_basic_live"] * S
    sampler_rate = sampler_bits / (prof.sampler_ms * 1e-3) if prof.sampler_ms > 0 else 0.0

    line = None
    if rank == 0:
        # Integer roofline of the sampler.  Measured on sm_100 (canopy_microbench_int): LOP3, IMAD.WIDE.U32 and
        # any mix of the two issue at ONE shared rate of 64 lanes/clk/SM, so the bound is the integer instruction count.
        int_peak, _ = microbench_int(0, local)
        kernel_name = {1: "sampler_direct_kernel", 2: "sampler_sliced_kernel", 4: "sampler_classed_kernel"}.get(spec, "sampler") + f"<{args.threshold_bits}>"
        traffic, exec_x_events, cap_note = captured(args.workload)
        exec_ops_per_bit = exec_x_events / stats["n_basic_live"] if exec_x_events and stats["n_basic_live"] else None
        step_ms = ms / args.steps
        step_bytes = (gate_bytes_per_sample + sampler_bytes_per_sample) * S
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": f"u{args.threshold_bits}", "data": "synthetic",
            "config": dict(config_block(args.workload, g, S, S_full, stats), threshold_bits=args.threshold_bits),
            "gate_evals_per_sec": value * stats["n_gates_live"],
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(r_e2e.h2d_bytes),
                    "d2h_bytes_per_step": int(nt * 8), "steps": e2e_steps,
                    "host_ms_per_step": {k2: v / e2e_steps for k2, v in e2e_parts.items()},
                    "what": "canopy_plan_create from host arrays + upload + run + allreduce + tallies to host, wall clock"},
            "e2e_cold": e2e_cold,
            "gpu_launches": launches_total,
            "collective": ("one all_reduce of the [steps, n_tallies] int64 tallies after the K timed steps, inside the timed region"
                           if world > 1 else "none (N = 1)"),
            "rank_ms": rank_ms,
            "roofline": {"kernel": "gate_ring_kernel" if prof.sampler_launches else "small_model_kernel (sampler + walk fused over shared-memory rows: the HBM figure is nominal, nothing but tallies leaves the SM)", "bound": "hbm", "achieved": gate_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": gate_gbs / peaks["hbm_gbs"], "traffic": traffic,
                         "traffic_note": "DRAM read+write bytes of one full launch, " + cap_note + "; "
                                         "frac can read a little above 1: the peak is a 1:1 read:write copy, this kernel's traffic is about 5:1 reads",
                         "alg_bytes_per_launch": gate_bytes_per_launch, "nominal_hbm_gbs": 7700.0, "frac_of_nominal": gate_gbs / 7700.0,
                         "peak_source": peaks_src, "alg_bytes_per_sample": gate_bytes_per_sample, "avg_launch_ms": gate_launch_ms,
                         "share_of_step": prof.gate_ms / max(prof.total_ms, 1e-9),
                         "timed": "profile pass, kernels serialised; `frac` = frac_alone: the walk launched as a kernel that owns the GPU",
                         "frac_alone": gate_gbs / peaks["hbm_gbs"],
                         "frac_step_shape": (gate_bytes_per_sample * S / (prof_shape.gate_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if prof_shape.gate_ms > 0 else None,
                         "avg_launch_ms_step_shape": prof_shape.gate_ms / max(1, prof_shape.gate_launches),
                         "step_shape": "one walk CTA per SM, as every timed step launches it beside the next batch's sampler (here: alone)",
                         "alg_bytes_full_launch": gate_bytes_per_sample * prof.batch_slabs * prof.slab_samples,
                         "traffic_over_alg_full_launch": (traffic / (gate_bytes_per_sample * prof.batch_slabs * prof.slab_samples)) if traffic else None},
            "roofline_sampler": {"kernel": kernel_name,
                                 "bound": "int (LOP3 + IMAD.WIDE share one issue rate of 64 lanes/clk/SM)",
                                 "achieved": sampler_rate * alg_ops_per_bit / 1e9, "peak": int_peak / 1e9, "unit": "G int lane-instr/s",
                                 "frac": sampler_rate * alg_ops_per_bit / int_peak,
                                 "executed_frac": sampler_rate * exec_ops_per_bit / int_peak if exec_ops_per_bit else None,
                                 "event_mix": sampler_mix,
                                 "bernoulli_bits_per_s": sampler_rate, "alg_int_ops_per_bit": alg_ops_per_bit,
                                 "executed_int_ops_per_bit": exec_ops_per_bit,
                                 "hbm_write_GBs": sampler_bytes_per_sample * S / (prof.sampler_ms * 1e-3) / 1e9 if prof.sampler_ms > 0 else 0.0,
                                 "avg_launch_ms": prof.sampler_ms / max(1, prof.sampler_launches),
                                 "share_of_step": prof.sampler_ms / max(prof.total_ms, 1e-9),
                                 "peak_source": "canopy_microbench_int(kind=0) measured in this run; executed counts: " + cap_note},
            # the pipelined step against its two resources: what the step would take if the two kernels overlapped perfectly
            # (max of the two, profile pass) and what the HBM alone allows (all bytes of the step at the measured copy peak)
            "roofline_step": {"ms_per_step": step_ms, "sampler_ms_alone": prof.sampler_ms, "gate_ms_alone": prof.gate_ms,
                              "frac_of_max_kernel": max(prof.sampler_ms, prof.gate_ms) / step_ms if step_ms > 0 else None,
                              "alg_bytes_per_sample": gate_bytes_per_sample + sampler_bytes_per_sample,
                              "achieved_GBs": step_bytes / (step_ms * 1e-3) / 1e9 if step_ms > 0 else None,
                              "frac_hbm": step_bytes / (step_ms * 1e-3) / 1e9 / peaks["hbm_gbs"] if step_ms > 0 else None},
            "tally_check": check if check is not None else "skipped (--no-check)",
            "top_event_tally_last_step": top_tally,
            "build_tag": build_tag(),
        }
    if world == 1 and rank == 0:
        cpu_v, cores, n_cpu, dt_cpu, _ = cpu_arm(g, S, tally_all, spec=spec, bits=args.threshold_bits)
        line["cpu_baseline"] = {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{n_cpu} samples of {args.workload} in {dt_cpu:.1f} s (bit-packed pthreads port in oracle/)"}
        if not args.no_study:
            line["other_configs"] = {n: side_line(n, local, stream, args.sampler, bits=args.threshold_bits) for n in ("c2", "c3") if n != args.workload}
            # the same workload under the other threshold width (64-bit thresholds are what p below 1e-8 needs), a quarter step
            ob = 96 - args.threshold_bits
            line["other_configs"][f"{args.workload}_u{ob}"] = side_line(args.workload, local, stream, args.sampler, steps=2, bits=ob, samples=max(S // 4, 1 << 20))
            if args.workload != "c1":   # config 1: the MEF example through the fused engine
                line["other_configs"]["c1"] = side_line("c1", local, stream, args.sampler, bits=args.threshold_bits)
            if spec != 2:   # the literal per-sample compare north_star words (spec 2, "sliced"), a quarter step
                line["other_configs"][f"{args.workload}_spec2"] = side_line(args.workload, local, stream, 2, steps=2, bits=args.threshold_bits, samples=max(S // 4, 1 << 20))
            if S_full != S and not args.samples:   # the configuration at its stated sample count, as ONE step on the driver's clock
                line["other_configs"][f"{args.workload}_full"] = side_line(args.workload, local, stream, args.sampler, steps=1, bits=args.threshold_bits, samples=S_full, warmups=0)
    if rank == 0:
        if args.samples:
            line["config"]["reduced"] = f"--samples {args.samples}: side study, NOT the workload's sample count"
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
