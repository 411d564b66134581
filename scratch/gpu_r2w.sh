#!/bin/bash
# variant check: bit-exact render tests with the variant library, then the same-box A/B
v=$1
cp jplspatial_b200/libjplg.so /tmp/prod.so
cp scratch/libjplg_$v.so jplspatial_b200/libjplg.so
python -m pytest tests/test_gpu_render.py tests/test_gpu_configs.py -x -q 2>&1 | tail -n 3
cp /tmp/prod.so jplspatial_b200/libjplg.so
bash scratch/ab_variants.sh prod $v
