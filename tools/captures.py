"""tools/captures.py TAG [cfg ...]: read gpurun_out/TAG_{gate,sampler}_CFG.ncu-rep (written by tools/gpu_round.sh with NCU=1) and
write profiles/captures.json -- per workload the DRAM bytes of one gate-walk launch and the executed integer instructions per
Bernoulli bit of one sampler launch, with the tag of the build they were measured on (bench.py quotes them only for that build).
Also writes the per-capture summaries profiles/TAG_ncu_full_{gate,sampler}_CFG.csv and the sampler opcode table."""
import collections, csv, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from canopy_b200 import synth

INT = ("LOP3", "IMAD", "ISETP", "IADD3", "VIADD", "SHF", "SEL", "LEA", "MOV", "PRMT", "IABS", "PLOP3", "POPC", "VIMNMX", "IADD", "LOP", "FLO", "BREV", "SGXT", "BMSK")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return dict(zip(rows[0], rows[2])), dict(zip(rows[0], rows[1]))


def opcodes(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if "Address" in r and "Instructions Executed" in r)
    h = rows[hi]
    si, ei, ti = h.index("Source"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
    ops, thr = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) > max(ei, ti) and r[ei].isdigit():
            op = re.sub(r"^@!?U?P\d+\s+", "", r[si].strip()).split()[0].rstrip(";")
            ops[op] += int(r[ei]); thr[op] += int(r[ti])
    return ops, thr


def fnum(s):
    return float(s.replace(",", ""))


def main():
    tag = sys.argv[1]
    cfgs = sys.argv[2:] or ["c4", "c2"]
    path = os.path.join(ROOT, "profiles", "captures.json")
    try:
        cur = json.load(open(path))
    except Exception:
        cur = {}
    for cfg in cfgs:
        g = synth.config_graph(cfg)
        entry = {"build_tag": bench.build_tag(), "capture": tag, "slabs_per_launch": 296}
        rep = os.path.join(ROOT, "gpurun_out", f"{tag}_gate_{cfg}.ncu-rep")
        if os.path.exists(rep):
            v, u = raw(rep)
            rd = fnum(v["dram__bytes_read.sum"]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u["dram__bytes_read.sum"]]
            wr = fnum(v["dram__bytes_write.sum"]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u["dram__bytes_write.sum"]]
            entry.update(gate_dram_bytes_per_launch=rd + wr, gate_dram_read=rd, gate_dram_write=wr,
                         gate_ncu_us=fnum(v["gpu__time_duration.sum"]) * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u["gpu__time_duration.sum"], 1))
            ops, thr = opcodes(rep)
            entry["gate_warp_instr_per_launch"] = sum(ops.values())
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep,
                            os.path.join(ROOT, "profiles", f"{tag}_ncu_full_gate_{cfg}.csv")], check=True, stdout=subprocess.DEVNULL)
        rep = os.path.join(ROOT, "gpurun_out", f"{tag}_sampler_{cfg}.ncu-rep")
        if os.path.exists(rep):
            ops, thr = opcodes(rep)
            iw = sum(n for op, n in ops.items() if op.split(".")[0] in INT)
            entry["sampler_warp_instr_per_launch"] = sum(ops.values())
            entry["sampler_int_warp_instr_per_launch"] = iw
            entry["sampler_active_lane_frac"] = sum(thr.values()) / max(1, 32 * sum(ops.values()))
            entry["sampler_executed_int_ops_per_bit_x_events"] = iw * 32.0 / (296 * 8192)   # divide by the plan's live events
            v, u = raw(rep)
            entry["sampler_ncu_us"] = fnum(v["gpu__time_duration.sum"]) * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u["gpu__time_duration.sum"], 1)
            with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_source_sampler_{cfg}_opcodes.csv"), "w") as f:
                f.write("opcode,warp_instr_executed,thread_instr_executed\n")
                for op, n in ops.most_common():
                    f.write(f"{op},{n},{thr[op]}\n")
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep,
                            os.path.join(ROOT, "profiles", f"{tag}_ncu_full_sampler_{cfg}.csv")], check=True, stdout=subprocess.DEVNULL)
        cur[cfg] = entry
        print(cfg, json.dumps(entry))
    json.dump(cur, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
