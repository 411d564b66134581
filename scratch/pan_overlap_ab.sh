#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_render.py -m gpu -x -q -p no:hypothesispytest > gpurun_out/pytest_render.log 2>&1; tail -3 gpurun_out/pytest_render.log
for opt in "" "--no-pan-overlap" "--no-pan-overlap --no-push-ahead"; do
tag=$(echo "x$opt" | tr -d ' -')
timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e $opt > gpurun_out/bench_ab_$tag.json 2>gpurun_out/bench_ab_$tag.err || tail -5 gpurun_out/bench_ab_$tag.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_ab_$tag.json"))
print("$tag", round(d["ms_per_step"],4), d["bus_crc32"], d["gpu_launches"], {k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms_per_step"].items()}, d["clocks"])
PY
done
