mkdir -p gpurun_out
python -m pytest tests/test_render.py tests/test_video.py -x -q 2>&1 | tail -4
python tools/bench_render.py 2>&1 | tail -1 | tee gpurun_out/bench_render.json
python tools/bench_render.py --frames 64 2>&1 | tail -1 | tee -a gpurun_out/bench_render.json
python tools/bench_render.py --frames 16 --iters 50 2>&1 | tail -1 | tee -a gpurun_out/bench_render.json
