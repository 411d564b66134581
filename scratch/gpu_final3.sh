#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02y3_gpu_tests.txt 2>&1; tail -n 3 gpurun_out/r02y3_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02y3_smoke.txt 2>&1; tail -n 2 gpurun_out/r02y3_smoke.txt
python bench.py > gpurun_out/r02y3_bench_default.json 2> gpurun_out/r02y3_bench_default.err; echo "bench rc=$?"
python bench.py --config 5 > gpurun_out/r02y3_bench_config5.json 2> gpurun_out/r02y3_bench_config5.err; echo "config 5 rc=$?"
python bench.py --config 4 > gpurun_out/r02y3_bench_config4.json 2> gpurun_out/r02y3_bench_config4.err; echo "config 4 rc=$?"
python - <<'PY'
import json
for f in ("default","config5","config4"):
    try:
        d=json.loads(open(f"gpurun_out/r02y3_bench_{f}.json").read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("frac"), d.get("gpu_launches"))
    except Exception as e: print(f, "ERR", e)
PY
