#!/bin/bash
# full validation of the tree: GPU suite, smoke, default bench (parity), throughput, reference arm
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -p no:hypothesispytest > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_parity.json 2> gpurun_out/bench_parity.err || tail -5 gpurun_out/bench_parity.err
timeout 600 python bench.py --mode throughput --no-cpu-baseline > gpurun_out/bench_throughput.json 2> gpurun_out/bench_throughput.err || tail -5 gpurun_out/bench_throughput.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err || tail -5 gpurun_out/bench_reference.err
python - <<PY
import json
for t in ("parity","throughput","reference"):
    d=json.load(open(f"gpurun_out/bench_{t}.json"))
    print(t, d.get("ms_per_step"), d.get("value"), (d.get("e2e") or {}).get("value"), d.get("bus_crc32"), d.get("gpu_launches"), d.get("clocks"))
PY
