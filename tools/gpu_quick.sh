mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "frame_stack" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
