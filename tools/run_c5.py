"""BASELINE.json config 5 in full: the 100k-event fault tree (config c4's DAG), 2^36 samples in total, sharded over the
ranks by Philox counter range, one NCCL all-reduce of the tallies.  Launch: torchrun --nproc-per-node N tools/run_c5.py
[--log2-total 36].  Device-timed (CUDA events, max over ranks); rank 0 prints one JSON line."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from canopy_b200 import synth
from canopy_b200.dist import run_sharded
from canopy_b200.engine import Plan, estimate

ap = argparse.ArgumentParser()
ap.add_argument("--log2-total", type=int, default=36)
args = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
g = synth.config_graph("c4")
plan = Plan(g)
total = 1 << args.log2_total
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
run_sharded(plan, world << 24, seed=0xC5, sample_begin=1 << 50, device=dev)  # warm-up: buffers, NCCL, clocks
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
t = run_sharded(plan, total, seed=0xC5, sample_begin=0, device=dev)
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    st = plan.stats
    sec = float(ms.item()) * 1e-3
    top = int(t[0].item())
    p_hat, sigma, lo, hi = estimate(top, total)
    print(json.dumps({"config": "c5: 100k-event fault tree, 2^%d samples over %d GPU(s)" % (args.log2_total, world), "n_gpus": world,
                      "samples": total, "seconds": sec, "samples_per_s": total / sec, "gate_evals_per_s": total / sec * st["n_gates_live"],
                      "B_live": st["n_basic_live"], "G_live": st["n_gates_live"], "E_live": st["n_edges_live"], "layers": st["n_layers"],
                      "top_event_tally": top, "p_hat": p_hat, "sigma": sigma, "ci95": [lo, hi]}), flush=True)
if world > 1:
    dist.destroy_process_group()
