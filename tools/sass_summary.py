"""tools/sass_summary.py OUT_PREFIX: static evidence of what the built library contains (cuobjdump -sass canopy_b200/libcanopy.so):
per kernel the opcode histogram, the register count, and every line that carries a TMA / mbarrier / warp-reduction /
bulk-store mnemonic (UTMALDG, UBLKCP, UTMASTG, SYNCS, REDUX, ELECT, RED, ATOMS).  Writes OUT_PREFIX_<kernel>.txt."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "canopy_b200", "libcanopy.so")
prefix = sys.argv[1]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
regs = dict(re.findall(r"Function (\S+?):\s*\n\s*REG:(\d+)", res))
want = {"gate_ring_kernel_tILb1": "gate_ring_g4", "sampler_classed_kernelILi32": "sampler_classed_32",
        "small_model_kernelILi32ELb1": "small_model_32_classed", "sampler_sliced_kernelILi32": "sampler_sliced_32"}
KEY = re.compile(r"UTMALDG|UBLKCP|UTMASTG|SYNCS|REDUX|ELECT|\bRED\b|ATOMS|ATOMG|UTMAPF|FENCE|MEMBAR")
for chunk in sass.split("Function : ")[1:]:
    name = chunk.split("\n", 1)[0].strip()
    tag = next((v for k, v in want.items() if k in name), None)
    if not tag:
        continue
    lines = [l for l in chunk.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l)]
    ops = collections.Counter()
    keyed = []
    for l in lines:
        body = re.sub(r"/\*[0-9a-f]+\*/", "", l, count=1).split("/*")[0].strip().rstrip(";").strip()
        body = re.sub(r"^@!?U?P\d+\s+", "", body)
        if not body:
            continue
        op = body.split()[0]
        ops[op.split(".")[0]] += 1
        if KEY.search(op):
            keyed.append(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).strip())
    with open(f"{prefix}_{tag}.txt", "w") as f:
        f.write(f"kernel   {name}\nlibrary  canopy_b200/libcanopy.so (sm_100a)   registers {regs.get(name, '?')}   SASS instructions {sum(ops.values())}\n\n")
        f.write("opcode histogram (static):\n")
        for op, n in ops.most_common():
            f.write(f"  {op:<12}{n}\n")
        f.write("\nTMA / mbarrier / reduction / atomic instructions (address, instruction):\n")
        for l in keyed:
            f.write("  " + l + "\n")
    print(tag, sum(ops.values()), "instructions,", len(keyed), "keyed lines")
