#!/bin/bash
# round 2, lease B: warp-arbitration experiments on the fused kernel + ncu of the split path's kernels
mkdir -p gpurun_out
echo "== quick parity (current tree)"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lockstep or golden or ragged or gae or probe" > gpurun_out/pytest_quick.log 2>&1; tail -3 gpurun_out/pytest_quick.log
echo "== fused variants"
for v in f_hi f_lo f_lo_d8 f_hi_s8; do USVGPU_LIB=build_variants/$v.so timeout 300 python tools/time_step.py usv-asmc-simple 65536 2>&1 | tail -1 | sed "s/^/$v: /"; done
echo "== ncu split path"
export USVGPU_STEP=split
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain_split.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/prof_split $SHORT > gpurun_out/ncu_split.log 2>&1
tail -3 gpurun_out/ncu_split.log
ls -la gpurun_out | tail -5
