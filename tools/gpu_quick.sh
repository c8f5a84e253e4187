mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline --steps 300 > gpurun_out/t.json 2>gpurun_out/t.err; python -c "
import json; d=json.loads(open('gpurun_out/t.json').read().strip().splitlines()[-1]); print('%.4g'%d['value'], d['e2e']['value'], d['roofline']['kernels_us'])"
