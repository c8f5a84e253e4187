#!/bin/bash
# round 2, lease E (2 GPUs): new API tests, full GPU suite, multi-GPU correctness, 2-rank bench
mkdir -p gpurun_out
echo "== sb3 seat / info / options tests"
timeout 900 python -m pytest tests/test_sb3_seat.py -m gpu -x -q > gpurun_out/pytest_sb3.log 2>&1; tail -15 gpurun_out/pytest_sb3.log
echo "== full GPU suite"
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log
bash tools/gpu_multi.sh 2
