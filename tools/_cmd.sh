#!/bin/bash
# scratch: taper A/B under the PDL pipeline
out=gpurun_out; mkdir -p $out
for i in 1 2; do
CANOPY_NO_TAPER=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-study > $out/pdl_taper_off_$i.json 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --no-study > $out/pdl_taper_on_$i.json 2>&1
done
CANOPY_NO_TAPER=1 python tools/quick_var.py > $out/pdl_var_notaper.txt 2>&1
for f in $out/pdl_taper_o*.json; do echo "== $f"; grep -o '"value": [0-9.e+]*, "unit": "samples/s", "n_gpus": [0-9]*' $f | head -2; grep -o '"e2e": {"value": [0-9.e+]*' $f; done
cat $out/pdl_var_notaper.txt
