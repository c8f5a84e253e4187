#!/bin/bash
# round 2, lease A: fused step kernel vs the split path -- parity first, then timings, bench, ncu
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
echo "== quick parity (fused default)"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lockstep or golden or ragged or sb3 or rollout" > gpurun_out/pytest_fused_quick.log 2>&1; tail -4 gpurun_out/pytest_fused_quick.log
echo "== per-kernel timings"
for mode in fused split; do for n in 65536 262144 1048576; do USVGPU_STEP=$mode timeout 300 python tools/time_step.py usv-asmc-simple $n 2>&1 | tail -1 | sed "s/^/$mode: /"; done; done
for v in d8s16 d4s16 d8s12; do USVGPU_LIB=build_variants/$v.so timeout 300 python tools/time_step.py usv-asmc-simple 65536 2>&1 | tail -1 | sed "s/^/$v: /"; done
echo "== full GPU suite (fused default)"
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log
echo "== split path parity"
USVGPU_STEP=split timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu_split.log 2>&1; tail -3 gpurun_out/pytest_gpu_split.log
echo "== bench as the driver runs it"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_20.json 2> gpurun_out/bench_20.err; echo "rc $?"; tail -c 2500 gpurun_out/bench_20.json; tail -5 gpurun_out/bench_20.err
USVGPU_STEP=split timeout 900 python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline > gpurun_out/bench_20_split.json 2> gpurun_out/bench_20_split.err; echo "rc $?"; cut -c 1-400 gpurun_out/bench_20_split.json
echo "== ncu"
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
$SHORT > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu1.log 2>&1
$SHORT > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'step_fused' -s 30 -c 3 -o gpurun_out/prof $SHORT > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
