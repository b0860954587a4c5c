"""tools/ncu_summary.py REP.ncu-rep OUT.csv: selected counters of one `ncu --set full` capture as metric,unit,value."""
import csv, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ("Kernel Name", "gpu__time_duration", "dram__bytes", "dram__throughput", "lts__t_bytes", "lts__t_sector", "launch__",
        "sm__throughput", "sm__inst_executed_pipe", "sm__pipe_", "sm__inst_issued", "sm__issue_active", "sm__warps_active",
        "smsp__inst_executed.sum", "smsp__average_warps_issue_stalled", "smsp__cycles_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared",
        "smsp__warps_eligible", "sm__cycles_elapsed", "gr__ctas_launched")
with open(out, "w", newline="") as f:
    w = csv.writer(f); w.writerow(["metric", "unit", "value"])
    for h, u, v in zip(hdr, units, vals):
        if any(k in h for k in keep) and ".min." not in h and ".max." not in h:
            w.writerow([h, u, v])
print("wrote", out)
