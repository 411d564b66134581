#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_render.py tests/test_gpu_configs.py tests/test_gpu_scan.py -q 2>&1 | tail -n 3
for v in 1 0 1 0; do
JPLG_PUSH_GENERIC=$v python bench.py --config 4 --steps 50 --warmup 5 > gpurun_out/r2x_c4_$v.json 2>gpurun_out/r2x_c4_$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2x_c4_$v.json").read().strip().splitlines()[-1])
print("generic push=$v: config 4 ms/block", round(d["ms_per_step"],4), json.dumps(d.get("kernel_ms_per_step_rank0"))[:300])
PY
done
