"""tools/sampler_opcounts.py REP.ncu-rep N_EVENTS N_BLOCKS_PER_EVENT OUT.csv [OUT.json SPEC]
Executed warp instructions per opcode of one sampler launch (source page of an `ncu --import-source on` capture),
per warp-row of 4096 samples; with OUT.json also the integer-class total per Bernoulli bit that bench.py quotes."""
import collections, csv, json, re, subprocess, sys
rep, n_ev, n_blk, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if "Address" in r and "Instructions Executed" in r)
h = rows[hi]
si, ei, ti = h.index("Source"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
ops, thr = collections.Counter(), collections.Counter()
for r in rows[hi + 1:]:
    if len(r) > max(ei, ti) and r[ei].isdigit():
        op = re.sub(r"^@!?U?P\d+\s+", "", r[si].strip()).split()[0].rstrip(";")
        ops[op] += int(r[ei]); thr[op] += int(r[ti])
units = n_ev * n_blk
with open(out, "w") as f:
    f.write("opcode,warp_instr_executed,per_warp_row_of_4096_samples,thread_instr_executed\n")
    for op, n in ops.most_common():
        f.write(f"{op},{n},{n / units:.3f},{thr[op]}\n")
INT = ("LOP3", "IMAD", "ISETP", "IADD3", "VIADD", "SHF", "SEL", "LEA", "MOV", "PRMT", "IABS", "PLOP3", "POPC", "VIMNMX")
iw = sum(n for op, n in ops.items() if op.split(".")[0] in INT)
it = sum(thr[op] for op in ops if op.split(".")[0] in INT)
res = dict(warp_instr_per_row=sum(ops.values()) / units, int_warp_instr_per_row=iw / units,
           executed_int_ops_per_bit=iw / units * 32 / 4096, active_lane_int_ops_per_bit=it / units / 4096,
           imad_wide_per_row=ops["IMAD.WIDE.U32"] / units, source=rep.split("/")[-1])
print(json.dumps(res))
if len(sys.argv) > 6:
    path, spec = sys.argv[5], sys.argv[6]
    try:
        cur = json.load(open(path))
    except Exception:
        cur = {}
    cur[spec] = res
    json.dump(cur, open(path, "w"), indent=1)
