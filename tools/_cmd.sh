#!/bin/bash
# scratch: co-residency sweep for 1 KB rows
out=gpurun_out/rows_sweep.txt; : > $out
for rc in 1 2; do for sm in 0 30000 38000 47000 60000; do
  CANOPY_RING_CTAS=$rc CANOPY_SAMPLER_SMEM=$sm python tools/quick_overlap.py >> $out 2>&1
done; done
cat $out
