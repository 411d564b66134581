#!/bin/bash
# same-box comparison of bench.py argument sets: ab_args.sh "args A" "args B" ...
mkdir -p gpurun_out
i=0
for a in "$@"; do
i=$((i+1))
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e $a > gpurun_out/aa_$i.json 2>gpurun_out/aa_$i.err || tail -3 gpurun_out/aa_$i.err
python - <<PY
import json
d=json.load(open("gpurun_out/aa_$i.json"))
print("[$a]", round(d["ms_per_step"],5), d["bus_crc32"], "render", round(d["roofline"]["avg_launch_ms"]*1e3,1), {k:round(x*1e3,1) for k,x in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
