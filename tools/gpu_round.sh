set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench.json
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2>&1
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3"
$SHORT > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu1.log 2>&1
$SHORT > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'dynamics_kernel|sensor_kernel' -s 20 -c 4 -o gpurun_out/prof $SHORT > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
