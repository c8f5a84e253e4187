mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "replay or frame_stack or collect" 2>&1 | tail -4
python tools/bench_replay.py --envs 4096 2>&1 | tail -1 | tee gpurun_out/bench_replay.jsonl
python tools/bench_replay.py --envs 65536 --buffer-size 1048576 --steps 100 2>&1 | tail -1 | tee -a gpurun_out/bench_replay.jsonl
python tools/bench_rollout.py 2>&1 | tail -4 | tee gpurun_out/bench_rollout.jsonl
