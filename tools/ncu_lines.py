"""tools/ncu_lines.py REP.ncu-rep CUBIN KERNEL_SUBSTR UNITS: executed warp instructions per source line.
Joins the SASS page of an `ncu --import-source on` capture with `nvdisasm --print-line-info` of the same cubin
(by instruction offset) and prints, per source line, executed warp instructions / UNITS, split by issue class."""
import csv, re, subprocess, sys, collections
rep, cubin, kern, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[1]; ai, si, ei = h.index("Address"), h.index("Source"), h.index("Instructions Executed")
ex = []
for r in rows[2:]:
    if len(r) > ei:
        ex.append((int(r[ai], 16), r[si].strip(), int(r[ei])))
a0 = ex[0][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith("\t.section\t.text.") and kern in l)
line = None; off2line = {}
for l in dis[start + 1:]:
    if l.startswith("\t.section"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: line = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: off2line[int(m.group(1), 16)] = line
def cls(op):
    op = op.split(".")[0]
    if op.startswith("U") and op not in ("UBLKCP",): return "uniform"
    if op in ("LOP3","IMAD","ISETP","IADD3","VIADD","SHF","SEL","LEA","MOV","PRMT","IABS","PLOP3","POPC","VIMNMX"): return "int"
    return "other"
agg = collections.defaultdict(lambda: collections.Counter())
for a, src, n in ex:
    op = re.sub(r"^@!?U?P\d+\s+", "", src).split()[0].rstrip(";")
    agg[off2line.get(a - a0)][cls(op)] += n
tot = collections.Counter()
out = []
for k, c in agg.items():
    t = sum(c.values()); tot.update(c)
    out.append((t, k, c))
out.sort(key=lambda x: -x[0])
print(f"total per unit: int {tot['int']/units:.2f} uniform {tot['uniform']/units:.2f} other {tot['other']/units:.2f}")
for t, k, c in out[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
    print(f"{t/units:7.2f}  int {c['int']/units:6.2f} uni {c['uniform']/units:6.2f} oth {c['other']/units:6.2f}  {k}")
