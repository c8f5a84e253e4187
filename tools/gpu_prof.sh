#!/bin/bash
# one ncu --set full pass over the two step kernels of the in-tree library (or USVGPU_LIB): tools/gpu_prof.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/prof_$TAG $SHORT > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log
