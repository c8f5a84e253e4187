#!/bin/bash
# round 2, lease C: leaner split kernels (32-bit env index, compacted geo/rec, raw cp.async, L2 prefetch of the episode constants)
mkdir -p gpurun_out
echo "== quick parity (current tree)"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lockstep or golden or ragged or sensors or probe" > gpurun_out/pytest_quick.log 2>&1; tail -3 gpurun_out/pytest_quick.log
echo "== timings"
for n in 65536 262144; do timeout 300 python tools/time_step.py usv-asmc-simple $n 2>&1 | tail -1; done
for v in dynu2; do USVGPU_LIB=build_variants/$v.so timeout 300 python tools/time_step.py usv-asmc-simple 65536 2>&1 | tail -1 | sed "s/^/$v: /"; done
echo "== bench (driver's flags)"
timeout 900 python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline > gpurun_out/bench_20c.json 2> gpurun_out/bench_20c.err; echo "rc $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_20c.json')); print(d['value'], d['ms_per_step'], d['roofline']['kernels_us'], d['timing']['python_loop']['value'], d['e2e']['value'])"
echo "== ncu split path"
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain_c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/prof_c $SHORT > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log
