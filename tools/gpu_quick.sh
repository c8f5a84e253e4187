python tools/time_step.py usv-asmc-simple 65536
for k in 0 10 40; do USVGPU_LIB=build_variants/libusvgpu_s$k.so python tools/time_step.py usv-asmc-simple 65536 | sed "s/^/SUBSTEPS=$k /"; done
python tools/time_step.py usv-asmc-simple 262144
