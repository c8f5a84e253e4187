mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "frame_stack or replay or collect" 2>&1 | tail -3
python tools/bench_replay.py --envs 65536 --buffer-size 1048576 --steps 100 2>&1 | tail -1 | tee gpurun_out/bench_replay2.jsonl
python tools/bench_replay.py --envs 65536 --buffer-size 1048576 --steps 100 --stack 1 2>&1 | tail -1 | tee -a gpurun_out/bench_replay2.jsonl
