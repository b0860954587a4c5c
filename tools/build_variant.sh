#!/bin/bash
# tools/build_variant.sh NAME "-DFOO=1 -DBAR=2": tuning build of libcanopy into gpurun_out/variants/libcanopy_NAME.so
set -e
cd "$(dirname "$0")/../canopy_b200/csrc"
name=$1; defs=$2
mkdir -p build/$name build/variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
for f in engine multi; do
  nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -DCANOPY_BUILD -DCANOPY_SHARED -I../../include -I. $defs -c $f.cu -o build/$name/$f.o &
done
wait
nvcc $ARCH -shared -o build/variants/libcanopy_$name.so build/$name/engine.o build/$name/multi.o build/lowering.o build/capi.o build/mef_model.o build/uncertainty.o -ldl -lpthread
echo built $name
