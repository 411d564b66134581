#!/bin/bash
mkdir -p gpurun_out
echo "== tests"; timeout 900 python -m pytest tests/test_gpu_reverb.py tests/test_zz_no_alloc.py tests/test_gpu_erbus.py -x -q -m gpu 2>&1 | tail -n 3
cp jplspatial_b200/libjplg.so /tmp/prod.so
for rep in 1 2; do for v in prod fdnnopf; do
if [ $v = prod ]; then cp /tmp/prod.so jplspatial_b200/libjplg.so; else cp scratch/libjplg_$v.so jplspatial_b200/libjplg.so; fi
python bench.py --config 5 --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('config5 $v', d['ms_per_step'])"
python bench.py --config 5 --c5-serial --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('config5 serial $v', d['ms_per_step'])"
done; done
cp /tmp/prod.so jplspatial_b200/libjplg.so
