#!/bin/bash
mkdir -p gpurun_out
python scratch/one_c5.py 6 > gpurun_out/one_c5.log 2>&1 || { tail -n 5 gpurun_out/one_c5.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:k_fdn -s 4 -c 1 -f -o gpurun_out/r02zz_k_fdn python scratch/one_c5.py 6 > gpurun_out/ncu_fdn.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_render -s 4 -c 1 -f -o gpurun_out/r02zz_k_render_er python scratch/one_c5.py 6 > gpurun_out/ncu_er.log 2>&1
tail -n 2 gpurun_out/one_c5.log gpurun_out/ncu_fdn.log gpurun_out/ncu_er.log; ls -la gpurun_out/r02zz_*.ncu-rep
