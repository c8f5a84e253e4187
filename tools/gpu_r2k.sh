#!/bin/bash
# round 2, lease K (8 GPUs): shards on their own devices vs the single-GPU run, then bench.py exactly as the driver launches it
mkdir -p gpurun_out
N=8
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tests/_multi_gpu_worker.py 8195 40 usv-asmc-simple 2>&1 | grep -E "ranks|MULTI_GPU_OK|MISMATCH" | tee gpurun_out/multi_gpu_worker_g$N.log
for rep in a b; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_g${N}_$rep.json 2> gpurun_out/bench_g${N}_$rep.err; echo "rc $?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_g${N}_$rep.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['config']['envs_total'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.4f'%d['ms_per_step'], 'blocks', ['%.3f'%x for x in d['timing']['block_ms_max_over_ranks']], 'rank0', ['%.3f'%x for x in d['timing']['block_ms_rank0']], 'loop %.4g'%d['timing']['python_loop']['value'], 'enqueue us/step %.1f'%d['timing']['host_enqueue_us_per_step'], {k:'%.4g'%v['value'] for k,v in d.get('extra',{}).items()})
PY
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_g$N.json 2>/dev/null; cut -c1-300 gpurun_out/bench_ref_g$N.json
