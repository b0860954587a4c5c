#!/bin/bash
# tools/build_variant_full.sh NAME "-DFOO=1": like build_variant.sh, but every translation unit is rebuilt with the
# definitions (for spec constants that the host lowering reads too).  Output: canopy_b200/csrc/build/variants/libcanopy_NAME.so
set -e
cd "$(dirname "$0")/../canopy_b200/csrc"
name=$1; defs=$2
mkdir -p build/$name build/variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
for f in engine multi; do
  nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -DCANOPY_BUILD -DCANOPY_SHARED -I../../include -I. $defs -c $f.cu -o build/$name/$f.o &
done
for f in lowering capi mef_model uncertainty; do
  g++ -O2 -std=c++17 -fPIC -fvisibility=hidden -DCANOPY_BUILD -DCANOPY_SHARED -I../../include -I. -I/usr/local/cuda/include $defs -c $f.cpp -o build/$name/$f.o &
done
wait
nvcc $ARCH -shared -o build/variants/libcanopy_$name.so build/$name/engine.o build/$name/multi.o build/$name/lowering.o build/$name/capi.o build/$name/mef_model.o build/$name/uncertainty.o -ldl -lpthread
echo built $name
