#!/bin/bash
mkdir -p gpurun_out
python scratch/one_render.py 6 > gpurun_out/one_render.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render -s 3 -c 1 -f -o gpurun_out/r02a_k_render python scratch/one_render.py 6 > gpurun_out/ncu_render.log 2>&1
tail -3 gpurun_out/one_render.log gpurun_out/ncu_render.log
