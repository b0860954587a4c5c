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
    assert d["config"]["workload"].startswith("c1:") and "B=50 G=16" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == len(os.sched_getaffinity(0)) and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", CANOPY_BENCH_REF_BUDGET_S="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, text=True, capture_output=True, timeout=120)
    assert r.returncode == 0 and not [l for l in r.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_loads_no_native_code_of_the_product():
    """The CPU arm must time oracle/ only: importing bench's reference path must not load libcanopy.so, and its stdlib
    reader of the c1 example must produce the graph the library's reader produces."""
    code = ("import sys, bench; sys.argv=['bench.py','--impl','reference','--workload','c1','--steps','1','--warmup','0'];"
            "import os; os.environ['CANOPY_BENCH_REF_BUDGET_S']='0.5'; bench.main();"
            "maps=open('/proc/self/maps').read(); assert 'libcanopy' not in maps, 'libcanopy.so loaded in the reference arm'; "
            "assert 'liboracle' in maps")
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT, timeout=120)
    import numpy as np
    import bench
    from canopy_b200.engine import load_mef
    a = bench._c1_graph_stdlib(os.path.join(ROOT, "examples", "c1_small_ft.xml"))
    b = load_mef(os.path.join(ROOT, "examples", "c1_small_ft.xml"))
    for f in ("p", "op", "child_off", "child", "roots"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f


def test_reference_arm_with_64_bit_thresholds():
    env = dict(os.environ, CANOPY_BENCH_REF_BUDGET_S="1")
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1",
                                   "--steps", "1", "--warmup", "0", "--threshold-bits", "64"], env=env, text=True, timeout=300)
    d = json.loads([l for l in out.splitlines() if l.startswith("{")][0])
    assert d["dtype"] == "u64" and d["config"]["threshold_bits"] == 64 and d["value"] > 0


def test_recorded_cross_spec_law_check_on_config_c5():
    """Two recorded 2^36-sample runs of config c5 on 8 GPUs sampled the SAME graph under different specs -- spec 2 (one
    compare per sample, round 1) and spec 4 (classed sparse, round 2) -- with disjoint random streams.  If both obey the
    same Bernoulli law per sample, the two top-event estimates differ by less than 3 sigma of their difference."""
    import json
    import math
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    a, b = (json.loads(open(os.path.join(root, "profiles", f)).read().strip().splitlines()[-1])
            for f in ("r01f_c5_full_n8.json", "r02g_c5_full_n8.json"))
    assert a["samples"] == b["samples"] == 1 << 36 and a["B_live"] == b["B_live"] and a["G_live"] == b["G_live"]
    sigma_diff = math.hypot(a["sigma"], b["sigma"])
    assert a["top_event_tally"] != b["top_event_tally"]                  # different streams, not a replay
    assert abs(a["p_hat"] - b["p_hat"]) < 3.0 * sigma_diff, (a["p_hat"], b["p_hat"], sigma_diff)
