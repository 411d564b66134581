#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_panner.py tests/test_gpu_configs.py tests/test_gpu_frontend.py tests/test_gpu_lut_variants.py -q > gpurun_out/r2o_tests.txt 2>&1; tail -n 4 gpurun_out/r2o_tests.txt
for lanes in 16 8 16 8; do
JPLG_MDAP_LANES=$lanes python bench.py --config 2 --steps 200 --warmup 20 > gpurun_out/r2o_c2_$lanes.json 2>gpurun_out/r2o_c2_$lanes.err
JPLG_MDAP_LANES=$lanes python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/r2o_c3_$lanes.json 2>gpurun_out/r2o_c3_$lanes.err
python - <<PY
import json
a=json.loads(open("gpurun_out/r2o_c2_$lanes.json").read().strip().splitlines()[-1]); b=json.loads(open("gpurun_out/r2o_c3_$lanes.json").read().strip().splitlines()[-1])
print("lanes $lanes: config 2", round(a["ms_per_step"]*1e3,2), "us; config 3 step", round(b["ms_per_step"],5), "ms", {k:round(v*1e3,1) for k,v in b["roofline"]["kernel_ms_per_step"].items()}, b.get("bus_crc32"))
PY
done
python scratch/scan_ab.py 1024 16384 > gpurun_out/r2o_scan_ab.txt 2>/dev/null; cat gpurun_out/r2o_scan_ab.txt
