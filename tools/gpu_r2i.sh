#!/bin/bash
# round 2, lease I: dynamics CTA size A/B; full GPU suite on the current tree
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 3 --repeats 9"
for v in default dynb64 dynb128 default dynb64; do
  if [ $v = default ]; then unset USVGPU_LIB; else export USVGPU_LIB=build_variants/$v.so; fi
  timeout 600 $B > gpurun_out/bench_i_$v.json 2> gpurun_out/bench_i_$v.err; python -c "
import json; d=json.load(open('gpurun_out/bench_i_$v.json')); print('$v', '%.4g'%d['value'], '%.2f us/step'%(1e3*d['ms_per_step']), d['roofline']['kernels_us'], d['e2e']['variants'])"
done
unset USVGPU_LIB
echo "== full GPU suite"
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
