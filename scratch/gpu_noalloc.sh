#!/bin/bash
mkdir -p gpurun_out
echo "== alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -n 5
echo "== whole gpu suite"; timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -n 5
