mkdir -p gpurun_out
python bench.py --env usv-simple --steps 300 --warmup 20 --cpu-budget-s 6 > gpurun_out/bench_usv-simple.json 2> gpurun_out/bench_usv-simple.err; tail -c 600 gpurun_out/bench_usv-simple.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_usv-simple.json').read().strip().splitlines()[-1]); r=d['roofline']
print('%.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], r['kernels_us'], 'frac %.3f'%r['frac'], 'cpu %.4g'%d['cpu_baseline']['value'])
PY
python bench.py --env usv-simple --impl reference --steps 20 --warmup 3 | cut -c1-200
