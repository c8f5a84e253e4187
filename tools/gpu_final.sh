#!/bin/bash
# round 2, final-tree lease: full GPU suite, smoke, the bench as the driver runs it, the reference arm, one ncu pass
# usage: tools/gpu_final.sh [launches|full]   (one profiler pass per lease)
MODE=${1:-full}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/final_pytest_gpu.log
tail -4 gpurun_out/final_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/final_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc $?"
python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err; echo "bench default rc $?"
python - <<'PY'
import json
for f in ("final_bench", "final_bench_default"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "value %.4g" % d["value"], "ms/step %.4f" % d["ms_per_step"], "e2e %.4g" % d["e2e"]["value"], d["roofline"].get("kernels_us"),
          "frac %.3f" % d["roofline"]["frac"], d["clocks"])
PY
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; cut -c1-400 gpurun_out/final_bench_reference.json
SHORT="python bench.py --steps 6 --warmup 3 --preroll 2 --no-cpu-baseline --e2e-steps 3 --no-extra --repeats 1"
if [ "$MODE" = launches ]; then
  $SHORT > gpurun_out/final_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv $SHORT > gpurun_out/final_ncu_launches.log 2>&1
else
  $SHORT > gpurun_out/final_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sensor_kernel|dynamics_kernel' -s 40 -c 4 -o gpurun_out/final_prof $SHORT > gpurun_out/final_ncu_full.log 2>&1
fi
tail -2 gpurun_out/final_ncu_*.log
