#!/bin/bash
# default build + every tuning variant through tools/pack_probe.py (sampler/gate ms on c2, 2^24 samples)
out=gpurun_out/$1; : > $out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 >> $out
echo default >> $out; python tools/pack_probe.py >> $out 2>&1
for f in canopy_b200/csrc/build/variants/*.so; do echo $(basename $f) >> $out; CANOPY_LIB=$PWD/$f python tools/pack_probe.py >> $out 2>&1; done
cat $out
