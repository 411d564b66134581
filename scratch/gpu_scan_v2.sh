#!/bin/bash
# scratch/libjplg_scanv2.so (scratch/scan_v2.patch applied) against the tree's library, same box
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/prod.so
cp scratch/libjplg_scanv2.so jplspatial_b200/libjplg.so
echo "== v2 tests (blocked scan)"; timeout 600 python -m pytest tests/test_gpu_scan.py -x -q -m gpu 2>&1 | tail -n 3
echo "== v2 tests (JPLG_SCAN_SEQ=1)"; JPLG_SCAN_SEQ=1 timeout 600 python -m pytest tests/test_gpu_scan.py -x -q -m gpu 2>&1 | tail -n 3
echo "== v2 scan_ab blocked"; timeout 600 python scratch/scan_ab.py 256 1024 2048 4096 2>&1 | tail -n 5
echo "== v2 scan_ab JPLG_SCAN_SEQ=1"; JPLG_SCAN_SEQ=1 timeout 600 python scratch/scan_ab.py 256 1024 2048 2>&1 | tail -n 4
cp /tmp/prod.so jplspatial_b200/libjplg.so
echo "== tree library scan_ab"; timeout 600 python scratch/scan_ab.py 256 1024 2048 4096 2>&1 | tail -n 5
