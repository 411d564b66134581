#!/bin/bash
# r02e (1 GPU): smoke, new shim / mixing / config tests, config 4 with per-layout contexts, default bench
mkdir -p gpurun_out
timeout 240 python __graft_entry__.py --smoke > gpurun_out/r2e_smoke.txt 2>&1; rc=$?; tail -2 gpurun_out/r2e_smoke.txt
if [ $rc -ne 0 ]; then echo "SMOKE FAILED rc=$rc"; exit 1; fi
timeout 900 python -m pytest tests/test_gpu_shim.py tests/test_gpu_render.py tests/test_gpu_configs.py -m gpu -x -q -p no:hypothesispytest --timeout=400 2>&1 | tail -15 | tee gpurun_out/r2e_tests.txt
timeout 300 python bench.py --config 4 --steps 50 --warmup 5 > gpurun_out/r2e_bench_c4_1gpu.json 2> gpurun_out/r2e_bench_c4_1gpu.err; echo "1gpu c4 rc=$?"; tail -3 gpurun_out/r2e_bench_c4_1gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_c4_1gpu.json'))
print('c4 1gpu ms', d['ms_per_step'], d.get('kernel_ms_per_step_rank0'), d.get('bus_check'))"
timeout 600 python bench.py > gpurun_out/r2e_bench_default.json 2> gpurun_out/r2e_bench_default.err; echo "default rc=$?"; tail -3 gpurun_out/r2e_bench_default.err
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_default.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d.get('e2e'), 'roofline', d['roofline'].get('frac'))"
