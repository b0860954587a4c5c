#!/bin/bash
# tools/sanitize_host.sh: AddressSanitizer + UndefinedBehaviorSanitizer over the HOST side of libcanopy -- the lowering, the
# hand-written XML reader and MEF model builder, the expression tape / uncertainty loop and the C API glue -- by running the CPU
# test files that exercise them against an instrumented build (the CUDA objects are linked as they are).  No GPU needed.
# Writes profiles/r03_sanitize_host.log.  (compute-sanitizer, the device-side counterpart, is closed on this GPU pool.)
set -u
cd "$(dirname "$0")/.."
make -C canopy_b200/csrc -s -j4 || exit 1
B=canopy_b200/csrc/build/asan; mkdir -p $B
CXX=/usr/bin/g++   # the distribution compiler: its libasan / libubsan are installed
FLAGS="-w -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined -std=c++17 -fPIC -fvisibility=hidden -DCANOPY_BUILD -DCANOPY_SHARED -Iinclude -Icanopy_b200/csrc -I/usr/local/cuda/include"
for f in lowering capi capi_testing mef_model uncertainty; do $CXX $FLAGS -c canopy_b200/csrc/$f.cpp -o $B/$f.o || exit 1; done
$CXX -shared -fsanitize=address,undefined -o $B/libcanopy_asan.so $B/*.o canopy_b200/csrc/build/engine.o canopy_b200/csrc/build/multi.o \
    -L/usr/local/cuda/lib64 -lcudart_static -lrt -ldl -lpthread || exit 1
LOG=profiles/r03_sanitize_host.log
{
  echo "# $(date -u +%FT%TZ) ASan+UBSan host build: $($CXX --version | head -1)"
  LD_PRELOAD="$($CXX -print-file-name=libasan.so) $($CXX -print-file-name=libubsan.so)" \
  ASAN_OPTIONS=detect_leaks=0:abort_on_error=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  CANOPY_LIB=$PWD/$B/libcanopy_asan.so \
  python -m pytest tests/test_lowering.py tests/test_mef_model.py tests/test_mef_and_dist.py tests/test_session.py tests/test_uncertainty.py \
      -q -m "not gpu" -p no:cacheprovider 2>&1 | tail -25
  echo "exit=${PIPESTATUS[0]}"
} | tee $LOG
