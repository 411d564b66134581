#!/bin/bash
# A/B of two library builds on one box: scratch/libjplg_old.so vs the tree's libjplg.so
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/new.so
for rep in 1 2; do
for v in old new; do
if [ $v = old ]; then cp scratch/libjplg_old.so jplspatial_b200/libjplg.so; else cp /tmp/new.so jplspatial_b200/libjplg.so; fi
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2>gpurun_out/ab_$v.err || tail -3 gpurun_out/ab_$v.err
python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v.json"))
print("$v", round(d["ms_per_step"],5), d["bus_crc32"], {k:round(x*1e3,1) for k,x in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
done
cp /tmp/new.so jplspatial_b200/libjplg.so
