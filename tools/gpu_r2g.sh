#!/bin/bash
# round 2, lease G: integer compares in the sub-step, two phase-B passes taken together -- parity, A/B by the bench, ncu
mkdir -p gpurun_out
echo "== parity (current tree)"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_g.log 2>&1; tail -3 gpurun_out/pytest_g.log
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 3 --repeats 9"
for v in default fpcmp_u2 fpcmp_u1 intcmp_u1 nojam default; do
  if [ $v = default ]; then unset USVGPU_LIB; else export USVGPU_LIB=build_variants/$v.so; fi
  timeout 600 $B > gpurun_out/bench_g_$v.json 2> gpurun_out/bench_g_$v.err; python -c "
import json; d=json.load(open('gpurun_out/bench_g_$v.json')); print('$v', '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['roofline']['kernels_us'])"
done
unset USVGPU_LIB
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain_g.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/prof_g $SHORT > gpurun_out/ncu_g.log 2>&1
tail -2 gpurun_out/ncu_g.log
