python tools/time_step.py usv-asmc-simple 65536
for c in 2 4; do USVGPU_LIB=build_variants/libusvgpu_sc$c.so python tools/time_step.py usv-asmc-simple 65536 | sed "s/^/SENS_MIN_CTAS=$c /"; done
