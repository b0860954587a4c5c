"""Print the keys of a bench.py JSON line that a round's notes quote (usage: python tools/bench_show.py FILE)."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "n_gpus", d["n_gpus"])
print("e2e_cold", d.get("e2e_cold"))
print("rank_ms", d.get("rank_ms"))
r = d["roofline"]
print({k: r.get(k) for k in ("frac", "frac_alone", "frac_step_shape", "avg_launch_ms", "avg_launch_ms_step_shape", "traffic",
                             "traffic_over_alg_full_launch")})
print("roofline_step", d["roofline_step"])
print("tally_check", d["tally_check"])
for k, v in d.get("other_configs", {}).items():
    print(k, v["samples_per_s"], "ms", v["ms_per_step"], "spec", v["spec_version"], "S", v["samples_per_step"])
print("clocks", d["clocks"])
