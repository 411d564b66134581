#!/bin/bash
# warm-cache ncu --set full of k_render (no cache flush between replays: the steady state of the block loop)
mkdir -p gpurun_out
python scratch/one_render.py 8 > gpurun_out/one_render.log 2>&1 && \
ncu --set full --clock-control none --cache-control none --import-source on -k regex:k_render -s 5 -c 1 -f -o gpurun_out/r02i_k_render_warm python scratch/one_render.py 8 > gpurun_out/ncu_render.log 2>&1
tail -n 2 gpurun_out/one_render.log gpurun_out/ncu_render.log
