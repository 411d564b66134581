#!/bin/bash
# same-box A/B of library builds: scratch/ab_variants.sh v1 v2 ... (scratch/libjplg_<v>.so; "prod" = the tree's library), two rounds
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/prod.so
for rep in 1 2; do
for v in "$@"; do
if [ $v = prod ]; then cp /tmp/prod.so jplspatial_b200/libjplg.so; else cp scratch/libjplg_$v.so jplspatial_b200/libjplg.so; fi
python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2>gpurun_out/ab_$v.err || tail -n 3 gpurun_out/ab_$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/ab_$v.json").read().strip().splitlines()[-1])
print("$v", "step", round(d["ms_per_step"],5), "render launch us", round(d["roofline"]["avg_launch_ms"]*1e3,1), {k:round(x*1e3,1) for k,x in d["roofline"]["kernel_ms_per_step"].items()}, d.get("bus_crc32"), "frac", round(d["roofline"]["frac"],3))
PY
done
done
cp /tmp/prod.so jplspatial_b200/libjplg.so
