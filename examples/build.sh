#!/bin/bash
# Builds the out-of-tree example worlds against the public headers (include/) and libmadrona_b200.so:
# what a user of the library does with their own world. sm_100a; -fmad=false like the library, so that
# a world's float arithmetic rounds like the reference's host build.
set -e
cd "$(dirname "$0")/.."
LIBDIR=$PWD/gpu_ecs_madrona_b200
for ex in simple_taskgraph/simple force_field/force_field churn/churn; do
  src=examples/$ex.cu
  out=examples/$(dirname $ex)/lib$(basename $ex)_world.so
  if [ ! -f "$out" ] || [ "$src" -nt "$out" ] || [ -n "$(find include -newer "$out" -type f | head -1)" ]; then
    nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Iinclude \
      -Xcompiler -fPIC -shared "$src" -o "$out" -L"$LIBDIR" -lmadrona_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../gpu_ecs_madrona_b200'
  fi
done
