#!/bin/bash
# round 2, lease F: grid-shape / unroll variants by the bench, then the bench exactly as the driver runs it
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 3 --repeats 9"
for v in default dynu2 sens4x8 sens4x6 default dynu2; do
  if [ $v = default ]; then unset USVGPU_LIB; else export USVGPU_LIB=build_variants/$v.so; fi
  timeout 600 $B > gpurun_out/bench_f_$v.json 2> gpurun_out/bench_f_$v.err; python -c "
import json; d=json.load(open('gpurun_out/bench_f_$v.json')); print('$v', '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['roofline']['kernels_us'])"
done
unset USVGPU_LIB
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_f_driver.json 2> gpurun_out/bench_f_driver.err; echo "rc $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_f_driver.json')); print(json.dumps({k:d[k] for k in ('value','ms_per_step','roofline','roofline_fp64','issue_slot_util','targets','episode_mix')}, indent=1)[:3500])"
