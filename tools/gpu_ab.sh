#!/bin/bash
# A/B of build_variants/<name>.so against the in-tree library: short parity, then the graph-timed step and the per-kernel times
# usage: tools/gpu_ab.sh name [name ...]      (build them first with tools/build_variants.sh)
mkdir -p gpurun_out
run() {  # $1 = label, USVGPU_LIB set by the caller
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lockstep or golden or ragged or sensors or tie" 2>&1 | tail -1 | sed "s/^/$1 parity: /"
  for rep in 1 2; do
    timeout 600 python bench.py --steps 200 --warmup 10 --no-extra --no-cpu-baseline --e2e-steps 3 > gpurun_out/ab_$1.json 2> gpurun_out/ab_$1.err || { echo "$1 bench failed"; tail -3 gpurun_out/ab_$1.err; }
    python - "$1" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/ab_{sys.argv[1]}.json").read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g" % d["value"], "us/step %.2f" % (d["ms_per_step"] * 1e3), {k.split("::")[-1]: round(v, 2) for k, v in d["roofline"]["kernels_us"].items()})
PY
  done
}
run base
for v in "$@"; do USVGPU_LIB=build_variants/$v.so run $v; done
