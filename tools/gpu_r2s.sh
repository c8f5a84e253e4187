#!/bin/bash
# r02s: overlapped step (usvgpu_set_overlap): parity, then A/B at 1 M / 262 144 / 65 536 envs on one GPU
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "overlap" 2>&1 | tail -3 | tee gpurun_out/r02s_pytest.log
timeout 900 python tools/time_overlap.py 1048576 0 75776,2,8 151552,2,8 151552,2,6 113664,3,4 151552,3,4 131072,2,8 65536,2,8 227328,2,12 2>&1 | tee gpurun_out/r02s_ab_1M.log
timeout 600 python tools/time_overlap.py 262144 0 37888,2,8 75776,2,8 65536,2,8 56832,3,4 2>&1 | tee gpurun_out/r02s_ab_256k.log
timeout 600 python tools/time_overlap.py 65536 0 16384,2,8 32768,2,8 18944,2,8 2>&1 | tee gpurun_out/r02s_ab_64k.log
