mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 300 --warmup 10 > gpurun_out/bench_g$N.json 2> gpurun_out/bench_g$N.err; echo "rc $?"
tail -c 600 gpurun_out/bench_g$N.err
python bench.py --impl reference --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ref_g$N.json 2>&1
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_g$N.json').read().strip().splitlines()[-1])
print(d['n_gpus'], '%.4g'%d['value'], '%.4g'%d['e2e']['value'], d['ms_per_step'], d['clocks'], d['episode_stats']['env_steps'])
PY
