#!/bin/bash
# round 2, lease H: per-CTA shared-memory env tickets in the sensor kernel (A/B against the fixed walk), parity, ncu
mkdir -p gpurun_out
echo "== parity (current tree)"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_sb3_seat.py -m gpu -x -q > gpurun_out/pytest_h.log 2>&1; tail -3 gpurun_out/pytest_h.log
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 3 --repeats 9"
for v in default static_walk default static_walk; do
  if [ $v = default ]; then unset USVGPU_LIB; else export USVGPU_LIB=build_variants/$v.so; fi
  timeout 600 $B > gpurun_out/bench_h_$v.json 2> gpurun_out/bench_h_$v.err; python -c "
import json; d=json.load(open('gpurun_out/bench_h_$v.json')); print('$v', '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['roofline']['kernels_us'])"
done
unset USVGPU_LIB
for n in 4096 262144 1048576; do timeout 600 $B --envs-per-gpu $n --batches 2 > gpurun_out/bench_h_n$n.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_h_n$n.json')); print($n, '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['roofline']['kernels_us'])"; done
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain_h.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/prof_h $SHORT > gpurun_out/ncu_h.log 2>&1
tail -2 gpurun_out/ncu_h.log
