#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan.py -x -q > gpurun_out/r2k_scan_tests.txt 2>&1; tail -n 6 gpurun_out/r2k_scan_tests.txt
python scratch/one_scan.py 6 1024 > gpurun_out/one_scan.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 200 --csv --log-file gpurun_out/r2k_scan_launches.csv python scratch/one_scan.py 6 1024 > gpurun_out/ncu_scan.log 2>&1
tail -n 2 gpurun_out/one_scan.log
timeout 600 python scratch/scan_ab.py 256 1024 2048 4096 > gpurun_out/r2k_scan_ab.txt 2> gpurun_out/r2k_scan_ab.err; cat gpurun_out/r2k_scan_ab.txt; tail -n 3 gpurun_out/r2k_scan_ab.err
