mkdir -p gpurun_out
for v in "" build/libusvgpu_mb5.so build/libusvgpu_mb6.so; do
  for n in 65536 1048576; do
    USVGPU_LIB=${v:-gym_usv_b200/libusvgpu.so} python bench.py --env usv-asmc-v0 --envs-per-gpu $n --steps 300 --no-cpu-baseline --e2e-steps 3 > gpurun_out/t.json 2>gpurun_out/t.err
    python -c "
import json; d=json.loads(open('gpurun_out/t.json').read().strip().splitlines()[-1]); print('$v', $n, '%.3g'%d['value'], d['roofline']['kernels_us'], '%.3f'%d['roofline']['frac'])"
  done
done
