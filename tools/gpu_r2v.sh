#!/bin/bash
# r02v: the FP64 diet as the default build: full GPU suite + smoke on it, then A/B against the previous arithmetic (build_variants/old.so)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02v_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02v_pytest_gpu.log
tail -4 gpurun_out/r02v_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02v_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/r02v_smoke.log
bash tools/gpu_r2u.sh "$@" 2>&1 | tee gpurun_out/r02v_ab.log
