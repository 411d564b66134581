#!/bin/bash
mkdir -p gpurun_out
python scratch/ncu_pan.py > gpurun_out/ncu_pan_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_mdap_3d -s 2 -c 1 -f -o gpurun_out/r02p_k_mdap python scratch/ncu_pan.py > gpurun_out/ncu_pan.log 2>&1
tail -n 1 gpurun_out/ncu_pan.log
