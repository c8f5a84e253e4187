#!/bin/bash
# r02ad: timing-only A/B (no parity run: for knobs that do not change arithmetic), one pass at 65 536 and 1 M envs
mkdir -p gpurun_out
run() {
  for n in 65536 1048576; do
    steps=200; [ $n -gt 100000 ] && steps=24
    timeout 300 python bench.py --envs-per-gpu $n --steps $steps --warmup 5 --repeats 7 --preroll 300 --no-extra --no-cpu-baseline --e2e-steps 3 > gpurun_out/v_$1.json 2> gpurun_out/v_$1.err || { echo "$1 bench failed"; tail -3 gpurun_out/v_$1.err; }
    python - "$1" $n <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/v_{sys.argv[1]}.json").read().strip().splitlines()[-1])
n = int(sys.argv[2])
print(sys.argv[1], n, "value %.4g" % d["value"], "us/step %.2f" % (d["ms_per_step"] * 1e3), "us per 65536 envs %.2f" % (d["ms_per_step"] * 1e3 * 65536 / n),
      {k.split("::")[-1]: round(v, 2) for k, v in d["roofline"]["kernels_us"].items()})
PY
  done
}
run base
for v in "$@"; do USVGPU_LIB=build_variants/$v.so run $v; done
