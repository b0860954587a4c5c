#!/bin/bash
# run tools/quick_gate.py with the default build and every tuning variant; output to gpurun_out/$1
out=gpurun_out/$1; shift
echo default > $out; python tools/quick_gate.py "$@" >> $out 2>&1
for f in canopy_b200/csrc/build/variants/*.so; do echo $(basename $f) >> $out; CANOPY_LIB=$PWD/$f python tools/quick_gate.py "$@" >> $out 2>&1; done
cat $out
