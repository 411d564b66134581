#!/bin/bash
# last session of round 2 on one GPU (library rebuilt after the packed FDN filter and the k_scan_mix staging): smoke, contract line, reference arm, configs 5 and 2, launch list
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02y2_smoke.txt 2>&1; tail -n 2 gpurun_out/r02y2_smoke.txt
python bench.py > gpurun_out/r02y2_bench_default.json 2> gpurun_out/r02y2_bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02y2_bench_reference.json 2> gpurun_out/r02y2_bench_reference.err; echo "ref rc=$?"
for c in 5 2; do python bench.py --config $c > gpurun_out/r02y2_bench_config$c.json 2> gpurun_out/r02y2_bench_config$c.err; echo "config $c rc=$?"; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02y2_launches_c5.csv python bench.py --config 5 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02y2_ncu_c5.log 2>&1; echo "ncu c5 rc=$?"
python - <<'PY'
import json
for f in ("default","reference","config5","config2"):
    try:
        d=json.loads(open(f"gpurun_out/r02y2_bench_{f}.json").read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("frac"), d.get("gpu_launches"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
