#!/bin/bash
mkdir -p gpurun_out
for k in k_scan_front k_scan_mix; do
ncu --set full --clock-control none --cache-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/r02l_$k python scratch/one_scan.py 6 1024 > gpurun_out/ncu_$k.log 2>&1
tail -n 1 gpurun_out/ncu_$k.log
done
