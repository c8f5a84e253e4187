mkdir -p gpurun_out
python __graft_entry__.py smoke 2>&1 | tail -3
for e in usv-asmc-v0 usv-asmc-ye-int-v0 usv-pid-v0; do
  python bench.py --env $e --envs-per-gpu 1048576 --steps 100 --warmup 5 --cpu-budget-s 5 2>gpurun_out/bench_$e.err | tee gpurun_out/bench_$e.json | cut -c1-400
done
