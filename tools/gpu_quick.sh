timeout 1700 python -m pytest tests -x -q -m gpu 2>&1 | tail -n 8
for v in ca_64_8 ca_32_16 ca_128_4 ca_64_6; do echo $v; USVGPU_LIB=build_variants/$v.so timeout 300 python tools/bench_ca.py --cpu-envs 0 2>&1 | cut -c 1-190; USVGPU_LIB=build_variants/$v.so timeout 300 python tools/bench_ca.py --cpu-envs 0 --envs 65536 2>&1 | cut -c 1-190; done
echo default; timeout 300 python tools/bench_ca.py --cpu-envs 0 2>&1 | cut -c 1-190; timeout 300 python tools/bench_ca.py --cpu-envs 0 --envs 65536 2>&1 | cut -c 1-190
