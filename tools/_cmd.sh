out=gpurun_out/ovl2.txt; : > $out
python - >> $out 2>&1 <<'PY'
import sys; sys.path.insert(0,'.')
from canopy_b200.engine import microbench_int
for k,n in enumerate(["lop3","imad_wide(+lop3)","philox_mix","popc","imadw:2lop3"]):
    ops,ms=microbench_int(k); print(n, f"{ops/1e12:.2f} T lane-instr/s", f"{ms:.3f} ms")
PY
python tools/quick_overlap.py >> $out 2>&1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 >> $out
cat $out
