#!/bin/bash
# same-box A/B of k_fdn: packed FP32x2 crossover (tree) against the scalar one (scratch/libjplg_fdnscalar.so), config 5, two rounds
mkdir -p gpurun_out
python -m pytest tests/test_gpu_reverb.py tests/test_gpu_erbus.py -x -q -m gpu 2>&1 | tail -n 3 > gpurun_out/fdn_tests.txt
cat gpurun_out/fdn_tests.txt
cp jplspatial_b200/libjplg.so /tmp/prod.so
for rep in 1 2; do
for v in prod fdnscalar; do
if [ $v = prod ]; then cp /tmp/prod.so jplspatial_b200/libjplg.so; else cp scratch/libjplg_$v.so jplspatial_b200/libjplg.so; fi
python bench.py --config 5 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/fdn_$v.json 2>gpurun_out/fdn_$v.err || tail -n 3 gpurun_out/fdn_$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/fdn_$v.json").read().strip().splitlines()[-1])
print("$v", "step ms", round(d["ms_per_step"],5), "value", d["value"], d.get("roofline",{}).get("kernel_ms_per_step"), d.get("bus_crc32"))
PY
done
done
cp /tmp/prod.so jplspatial_b200/libjplg.so
