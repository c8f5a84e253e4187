#!/bin/bash
# host-face checks: the tests that go through usvgpu_step_host, then the e2e figures with and without the in-library graphs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "host or sb3 or seat or lazy or info or ragged or video or frame or replay" 2>&1 | tail -4
for rep in 1 2; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/e2e_$rep.json 2> gpurun_out/e2e_$rep.err || tail -5 gpurun_out/e2e_$rep.err
python - $rep <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/e2e_{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], {k: "%.4g" % v for k, v in d["e2e"]["variants"].items()})
PY
done
timeout 300 python tools/e2e_breakdown.py 2>&1 | tail -7
