#!/bin/bash
mkdir -p gpurun_out
echo "== whole gpu suite"; timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -n 4
echo "== scan_ab"; timeout 600 python scratch/scan_ab.py 256 1024 2048 4096 16384 2>&1 | tail -n 6
