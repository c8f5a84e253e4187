mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "gae or collect_rollouts" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
python tools/bench_rollout.py --envs 4096 --n-steps 2048 | tee gpurun_out/rollout_4096.json
python tools/bench_rollout.py --envs 65536 --n-steps 128 | tee gpurun_out/rollout_65536.json
python tools/bench_rollout.py --envs 65536 --n-steps 128 --env usv-asmc-simple | tee gpurun_out/rollout_65536_asmc.json
