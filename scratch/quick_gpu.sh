#!/bin/bash
# quick GPU check: parity tests + short bench in both modes (prints kernel times)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -p no:hypothesispytest 2>&1 | tail -5
for mode in parity throughput; do
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --mode $mode ${1---no-e2e} > gpurun_out/bench_q_$mode.json 2>gpurun_out/bench_q_$mode.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_q_$mode.json"))
print("$mode", round(d["ms_per_step"],4), {k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms_per_step"].items()}, (d.get("e2e") or {}).get("ms_per_step"))
PY
done
