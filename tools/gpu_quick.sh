mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
for i in 1 2; do python bench.py --no-cpu-baseline --steps 600 > gpurun_out/t.json 2>gpurun_out/t.err; python -c "
import json; d=json.loads(open('gpurun_out/t.json').read().strip().splitlines()[-1]); print('%.4g'%d['value'], '%.4g'%d['e2e']['value'], d['roofline']['kernels_us'])"; done
python bench.py --no-cpu-baseline --steps 300 --obstacles-f32 0 > gpurun_out/t64.json 2>gpurun_out/t.err; python -c "
import json; d=json.loads(open('gpurun_out/t64.json').read().strip().splitlines()[-1]); print('f64 table', '%.4g'%d['value'], d['roofline']['kernels_us'])"
