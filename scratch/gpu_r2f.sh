#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scratch/c4_serial.py 8192 2>&1 | tee gpurun_out/r2f_c4_serial.txt
timeout 300 python bench.py --config 4 --steps 50 --warmup 5 > gpurun_out/r2f_bench_c4_1gpu.json 2> gpurun_out/r2f_bench_c4_1gpu.err; echo "1gpu c4 rc=$?"; tail -3 gpurun_out/r2f_bench_c4_1gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r2f_bench_c4_1gpu.json'))
print('c4 1gpu ms', d['ms_per_step'], d.get('kernel_ms_per_step_rank0'), d.get('bus_check'))"
