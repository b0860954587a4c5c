#!/bin/bash
# tools/run_probe_variants.sh OUT [cfgs]: tools/probe_step.py with the default build and every tuning variant under build/variants
out=gpurun_out/$1; shift
mkdir -p gpurun_out
python tools/probe_step.py "$@" > $out 2>&1
for f in canopy_b200/csrc/build/variants/*.so; do CANOPY_LIB=$PWD/$f timeout 600 python tools/probe_step.py "$@" >> $out 2>&1; done
cat $out
