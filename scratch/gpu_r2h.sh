#!/bin/bash
# role accounting (PROF build) + ncu --set full of the current k_render
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/prod.so
cp scratch/libjplg_prof.so jplspatial_b200/libjplg.so
python scratch/render_prof.py > gpurun_out/r2h_prof.txt 2>&1
cp /tmp/prod.so jplspatial_b200/libjplg.so
cat gpurun_out/r2h_prof.txt
python scratch/one_render.py 6 > gpurun_out/one_render.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render -s 3 -c 1 -f -o gpurun_out/r02h_k_render python scratch/one_render.py 6 > gpurun_out/ncu_render.log 2>&1
tail -2 gpurun_out/one_render.log gpurun_out/ncu_render.log
