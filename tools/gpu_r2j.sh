#!/bin/bash
# round 2, lease J: host-face tests after the lazy infos change; e2e figures
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_sb3_seat.py tests/test_ca_gpu.py -m gpu -x -q -k "host or sb3 or seat or experiment or semantics or legacy" > gpurun_out/pytest_j.log 2>&1; tail -3 gpurun_out/pytest_j.log
B="python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline --e2e-steps 40 --repeats 3"
timeout 600 $B > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; python -c "
import json; d=json.load(open('gpurun_out/bench_j.json')); print('%.4g'%d['value'], d['e2e']['variants'])"
python tools/e2e_breakdown.py 2>&1 | tail -12
