#!/bin/bash
# multi-GPU lease (gpurun --gpus N): hardware correctness of the sharded path + the bench as the driver launches it
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader
echo "== multi-GPU correctness (torchrun + NCCL, shards on their own devices vs the single-GPU run)"
timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -rA > gpurun_out/pytest_multi_gpu.log 2>&1; tail -6 gpurun_out/pytest_multi_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tests/_multi_gpu_worker.py 4099 60 usv-asmc-simple 2>&1 | grep -E "ranks|MULTI_GPU_OK|MISMATCH" | tee gpurun_out/multi_gpu_worker_g$N.log
echo "== bench, $N ranks, the driver's flags"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_g$N.json 2> gpurun_out/bench_g$N.err; echo "rc $?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_g$N.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['config']['envs_total'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.4f'%d['ms_per_step'], 'blocks', d['timing']['block_ms_max_over_ranks'], 'loop %.4g'%d['timing']['python_loop']['value'], {k:'%.4g'%v['value'] for k,v in d.get('extra',{}).items()})
PY
