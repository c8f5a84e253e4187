#!/bin/bash
# round 2, lease O: programmatic dependent launch extended to the kinematic / legacy / CA step kernels -- full GPU suite on
# the in-tree library, then the side envs with the in-tree library against build_variants/old.so (the previous commit)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_o.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_o.log; tail -4 gpurun_out/pytest_o.log
one() {  # $1 label, rest: bench flags
  lab=$1; shift
  timeout 600 python bench.py --steps 200 --warmup 10 --no-extra --no-cpu-baseline --e2e-steps 3 "$@" > gpurun_out/o_$lab.json 2> gpurun_out/o_$lab.err || { echo "$lab failed"; tail -3 gpurun_out/o_$lab.err; }
  python - "$lab" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/o_{sys.argv[1]}.json").read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g" % d["value"], "us/step %.2f" % (d["ms_per_step"] * 1e3), "loop %.4g" % d["timing"]["python_loop"]["value"])
PY
}
for lib in new old; do
  if [ $lib = old ]; then export USVGPU_LIB=build_variants/old.so; fi
  one ${lib}_simple --env usv-simple
  one ${lib}_v0_1m --env usv-asmc-v0 --envs-per-gpu 1048576
  one ${lib}_v0_4k --env usv-asmc-v0 --envs-per-gpu 4096
  one ${lib}_pid_1m --env usv-pid-v0 --envs-per-gpu 1048576
  timeout 300 python tools/bench_ca.py --cpu-envs 0 2>&1 | cut -c1-200 | sed "s/^/${lib}_ca /"
done
