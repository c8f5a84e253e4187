mkdir -p gpurun_out
python -m pytest tests/test_replay.py -x -q 2>&1 | tail -8
python tools/bench_replay.py --envs 4096 2>&1 | tail -1 | tee gpurun_out/bench_replay.jsonl
python tools/bench_replay.py --envs 65536 --buffer-size 1048576 --steps 100 2>&1 | tail -1 | tee -a gpurun_out/bench_replay.jsonl
