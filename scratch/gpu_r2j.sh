#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan.py -x -q > gpurun_out/r2j_scan_tests.txt 2>&1; tail -n 15 gpurun_out/r2j_scan_tests.txt
timeout 600 python scratch/scan_ab.py > gpurun_out/r2j_scan_ab.txt 2> gpurun_out/r2j_scan_ab.err; cat gpurun_out/r2j_scan_ab.txt; tail -n 3 gpurun_out/r2j_scan_ab.err
