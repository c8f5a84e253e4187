mkdir -p gpurun_out
N=${1:-2}
python -m pytest tests -m gpu -x -q -k "sb3_semantics" 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 300 --warmup 10 > gpurun_out/bench_g$N.json 2> gpurun_out/bench_g$N.err; echo "rc $?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 300 --warmup 10 --envs-per-gpu $((1048576 / N)) --no-cpu-baseline > gpurun_out/bench_1m_g$N.json 2>> gpurun_out/bench_g$N.err; echo "rc $?"
python - <<PY
import json
for f in ('gpurun_out/bench_g$N.json','gpurun_out/bench_1m_g$N.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(d['n_gpus'], d['config']['envs_total'], '%.4g'%d['value'], '%.4g'%d['e2e']['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['episode_stats']['env_steps'])
PY
