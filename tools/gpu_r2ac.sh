#!/bin/bash
# r02ac (8 GPUs, final tree): bench.py as the driver launches it, then the same with 1 048 576 envs per GPU
mkdir -p gpurun_out
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
show() {
python - "$1" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d['n_gpus'], d['config']['envs_total'], 'value %.4g' % d['value'], 'e2e %.4g' % d['e2e']['value'], 'ms/step %.4f' % d['ms_per_step'],
      'blocks', ['%.3f' % x for x in d['timing']['block_ms_max_over_ranks']], 'loop %.4g' % d['timing']['python_loop']['value'],
      {k: '%.4g' % v['value'] for k, v in d.get('extra', {}).items()})
PY
}
$TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02ac_bench_g$N.json 2> gpurun_out/r02ac_bench_g$N.err; echo "rc $?"; show gpurun_out/r02ac_bench_g$N.json
$TR --master-port 29512 bench.py --gpus $N --steps 12 --warmup 3 --envs-per-gpu 1048576 --preroll 300 --no-extra --no-cpu-baseline --e2e-steps 3 > gpurun_out/r02ac_bench_g${N}_1M_per_gpu.json 2> gpurun_out/r02ac_bench_g${N}_1M_per_gpu.err; echo "rc $?"; show gpurun_out/r02ac_bench_g${N}_1M_per_gpu.json
