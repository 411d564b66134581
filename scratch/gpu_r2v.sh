#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scan.py -x -q > gpurun_out/r2v_scan_tests.txt 2>&1; tail -n 4 gpurun_out/r2v_scan_tests.txt
timeout 300 python scratch/scan_ab.py 256 1024 2048 4096 > gpurun_out/r2v_scan_ab_pipe.txt 2>gpurun_out/r2v_err.txt; cat gpurun_out/r2v_scan_ab_pipe.txt | cut -c1-400
JPLG_SCAN_PHASES=1 timeout 300 python scratch/scan_ab.py 1024 > gpurun_out/r2v_scan_ab_phases.txt 2>/dev/null; cat gpurun_out/r2v_scan_ab_phases.txt | cut -c1-400
python scratch/one_scan.py 6 1024 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 200 --csv --log-file gpurun_out/r2v_scan_launches.csv python scratch/one_scan.py 6 1024 > /dev/null 2>&1
