#!/bin/bash
# r02t: the re-created container's build of the final tree: full GPU suite, smoke, the bench as the driver runs it
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02t_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02t_pytest_gpu.log
tail -4 gpurun_out/r02t_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02t_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/r02t_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02t_bench.json 2> gpurun_out/r02t_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02t_bench.json").read().strip().splitlines()[-1])
print("value %.4g" % d["value"], "ms/step %.4f" % d["ms_per_step"], "e2e %.4g" % d["e2e"]["value"], d["roofline"].get("kernels_us"),
      "frac %.3f" % d["roofline"]["frac"], d["clocks"], {k: "%.4g" % v["value"] for k, v in d.get("extra", {}).items()})
PY
