#!/bin/bash
# A/B of an environment knob on one box: ab_env.sh NAME=VALUE [bench args]
mkdir -p gpurun_out
knob=$1; shift
for rep in 1 2; do
for v in off on; do
if [ $v = on ]; then export $knob; else unset ${knob%%=*}; fi
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e "$@" > gpurun_out/ab_$v.json 2>gpurun_out/ab_$v.err || tail -3 gpurun_out/ab_$v.err
python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v.json"))
print("$v", round(d["ms_per_step"],5), d["bus_crc32"], round(d["roofline"]["avg_launch_ms"]*1e3,1), {k:round(x*1e3,1) for k,x in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
done
