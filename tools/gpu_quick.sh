mkdir -p gpurun_out
python __graft_entry__.py smoke 2>&1 | tail -3
bash tools/gpu_round.sh > gpurun_out/round.log 2>&1; tail -5 gpurun_out/round.log
