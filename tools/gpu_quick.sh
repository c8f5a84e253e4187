mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "host_face or sb3" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python tools/e2e_breakdown.py 2>&1 | tee gpurun_out/e2e_breakdown.txt
