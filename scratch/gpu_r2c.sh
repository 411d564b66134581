#!/bin/bash
mkdir -p gpurun_out
timeout 240 python __graft_entry__.py --smoke > gpurun_out/r2c_smoke.txt 2>&1; rc=$?; tail -3 gpurun_out/r2c_smoke.txt
if [ $rc -ne 0 ]; then echo "SMOKE FAILED rc=$rc"; exit 1; fi
timeout 900 python -m pytest tests -m gpu -q -p no:hypothesispytest --timeout=240 2>&1 | tail -15 | tee gpurun_out/r2c_tests.txt
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], round(d["ms_per_step"], 5), d.get("bus_crc32"), round(d["roofline"]["avg_launch_ms"] * 1e3, 1), {k: round(x * 1e3, 1) for k, x in d["roofline"]["kernel_ms_per_step"].items()}, "frac", round(d["roofline"]["frac"], 3))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
Q="--steps 300 --warmup 20 --no-cpu-baseline --no-e2e --no-capacity --no-offline"
for sched in realtime packed; do
  timeout 200 python bench.py $Q --schedule $sched > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "new/$sched" gpurun_out/ab.json
done 2>&1 | tee gpurun_out/r2c_ab.txt
timeout 300 python bench.py --config 4 --steps 50 --warmup 5 > gpurun_out/r2c_bench_c4.json 2> gpurun_out/r2c_bench_c4.err; echo "config 4 rc=$?"; tail -2 gpurun_out/r2c_bench_c4.err; cut -c1-400 gpurun_out/r2c_bench_c4.json
