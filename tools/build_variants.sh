#!/bin/bash
# Alternative builds of libusvgpu.so for A/B runs on the GPU box (USVGPU_LIB=build_variants/<name>.so python ...).
# usage: tools/build_variants.sh name "-DFLAG=... -DFLAG=..." [name flags ...]
set -e
cd "$(dirname "$0")/.."
mkdir -p build_variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -Xptxas -v $flags \
       -o build_variants/$name.so gym_usv_b200/csrc/usvgpu.cu 2> build_variants/$name.log
  echo "== $name ($flags)"; grep -A2 -E "sensor_kernelILb1EfEE|dynamics_kernelILb0EE" build_variants/$name.log | grep -E "registers|spill" || true
done
