"""bench.py's reference arm (the CPU port on the host cores) prints the contract's JSON line; CPU only, a few seconds.
The GPU arm's line is checked by the driver on the GPU box; its keys are produced by the same code path (`config_block`)."""
import json
import os
import subprocess
import sys

from helpers import ROOT


def test_reference_arm_json_line():
    env = dict(os.environ, CANOPY_BENCH_REF_BUDGET_S="2")
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1",
                                   "--steps", "2", "--warmup", "1"], env=env, text=True, timeout=300)
    lines = [l for l in out.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "mc_samples_per_sec" and d["unit"] == "samples/s"
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1 and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["config"]["workload"].startswith("c1:") and d["config"]["spec_version"] == 3
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == len(os.sched_getaffinity(0)) and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", CANOPY_BENCH_REF_BUDGET_S="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, text=True, capture_output=True, timeout=120)
    assert r.returncode == 0 and not [l for l in r.stdout.splitlines() if l.startswith("{")]
