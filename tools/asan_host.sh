#!/bin/bash
# The host-side sources (BVH / oct-tree / camera builders, traversal-layout builder, OBJ import, PNG) built with
# -fsanitize=address,undefined and the CPU suite run against that library (no GPU needed; the CUDA objects are linked as built).
#   bash tools/asan_host.sh        -> prints the number of sanitizer reports and the suite's result
set -e
cd "$(dirname "$0")/.."
python -c "import importlib; importlib.import_module('gpu-raytracer_b200.build').build_library()" > /dev/null 2>&1
mkdir -p tools/_asan
CX=/usr/bin/g++
for f in okrt_host okrt_accel okrt_obj okrt_image; do
  $CX -std=c++17 -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -ffp-contract=off -fno-fast-math -fPIC -pthread \
      -fvisibility=hidden -c gpu-raytracer_b200/csrc/$f.cpp -o tools/_asan/$f.o -Iinclude -Igpu-raytracer_b200/csrc
done
/usr/local/cuda/bin/nvcc -ccbin $CX -shared -o tools/_asan/libokrt_asan.so tools/_asan/okrt_host.o tools/_asan/okrt_accel.o \
    tools/_asan/okrt_obj.o tools/_asan/okrt_image.o gpu-raytracer_b200/_obj/okrt_device.cu.o gpu-raytracer_b200/_obj/okrt_build_device.cu.o \
    gpu-raytracer_b200/_obj/okrt_layout_device.cu.o -lpthread -Xcompiler -fsanitize=address,-fsanitize=undefined 2>/dev/null
LD_PRELOAD="$($CX -print-file-name=libasan.so) $($CX -print-file-name=libubsan.so)" ASAN_OPTIONS=detect_leaks=0:halt_on_error=0 \
  UBSAN_OPTIONS=print_stacktrace=1 OKRT_LIB=$PWD/tools/_asan/libokrt_asan.so \
  python -m pytest tests -x -q -m "not gpu" -p no:cacheprovider > tools/_asan/run.log 2>&1 || true
echo "sanitizer reports: $(grep -c 'runtime error\|ERROR: AddressSanitizer' tools/_asan/run.log)"
tail -1 tools/_asan/run.log
