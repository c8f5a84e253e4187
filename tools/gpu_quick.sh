mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "legacy or sibling" 2>&1 | tail -3
for e in usv-asmc-v0 usv-asmc-ye-int-v0 usv-pid-v0; do
  python bench.py --env $e --envs-per-gpu 1048576 --steps 100 --warmup 5 --cpu-budget-s 5 2>gpurun_out/bench_$e.err | tee gpurun_out/bench_$e.json | cut -c1-200
done
python bench.py --env usv-asmc-v0 --envs-per-gpu 4096 --steps 300 --warmup 10 --no-cpu-baseline 2>/dev/null | tee gpurun_out/bench_legacy_4096.json | cut -c1-200
