#!/bin/bash
# A/B: scratch/libjplg_rows8.so (8-row MDAP batches) vs the tree's library (24-row)
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/prod.so
for rep in 1 2; do
for v in prod rows8; do
if [ $v = prod ]; then cp /tmp/prod.so jplspatial_b200/libjplg.so; else cp scratch/libjplg_$v.so jplspatial_b200/libjplg.so; fi
python bench.py --config 2 --steps 200 --warmup 20 > gpurun_out/r2q_c2_$v.json 2>gpurun_out/r2q_c2_$v.err
python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/r2q_c3_$v.json 2>gpurun_out/r2q_c3_$v.err
python - <<PY
import json
a=json.loads(open("gpurun_out/r2q_c2_$v.json").read().strip().splitlines()[-1]); b=json.loads(open("gpurun_out/r2q_c3_$v.json").read().strip().splitlines()[-1])
print("$v: config 2", round(a["ms_per_step"]*1e3,2), "us; config 3 step", round(b["ms_per_step"],5), "ms", {k:round(x*1e3,1) for k,x in b["roofline"]["kernel_ms_per_step"].items()}, b.get("bus_crc32"))
PY
done
done
cp /tmp/prod.so jplspatial_b200/libjplg.so
