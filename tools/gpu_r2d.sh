#!/bin/bash
# round 2, lease D: cp.async stash of the tail values, programmatic dependent launch (A/B against -DUSV_PDL=0)
mkdir -p gpurun_out
echo "== quick parity (current tree)"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lockstep or golden or ragged or sensors or host or sharding" > gpurun_out/pytest_quick.log 2>&1; tail -3 gpurun_out/pytest_quick.log
echo "== timings"
timeout 300 python tools/time_step.py usv-asmc-simple 65536 2>&1 | tail -1
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 3"
for v in default nopdl dynu2; do
  if [ $v = default ]; then unset USVGPU_LIB; else export USVGPU_LIB=build_variants/$v.so; fi
  timeout 600 $B > gpurun_out/bench_d_$v.json 2> gpurun_out/bench_d_$v.err; echo "$v rc $?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_d_$v.json')); print('$v', '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['timing']['block_ms_max_over_ranks'], 'loop %.4g'%d['timing']['python_loop']['value'])"
done
unset USVGPU_LIB
for n in 4096 262144; do timeout 600 $B --envs-per-gpu $n --batches 3 > gpurun_out/bench_d_n$n.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_d_n$n.json')); print($n, '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']))"; done
