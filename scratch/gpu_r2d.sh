#!/bin/bash
# r02d (2 GPUs): exchange tests, config 3 / config 4 at N = 2 with bus_check, issue-overlap microbenchmark, config 4 kernel breakdown at N = 1
mkdir -p gpurun_out
timeout 240 python __graft_entry__.py --smoke > gpurun_out/r2d_smoke.txt 2>&1; rc=$?; tail -2 gpurun_out/r2d_smoke.txt
if [ $rc -ne 0 ]; then echo "SMOKE FAILED rc=$rc"; exit 1; fi
timeout 600 python -m pytest tests/test_gpu_exchange.py -q -p no:hypothesispytest --timeout=400 2>&1 | tail -15 | tee gpurun_out/r2d_exchange_test.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/r2d_bench_2gpu.json 2> gpurun_out/r2d_bench_2gpu.err; echo "2gpu c3 rc=$?"; tail -3 gpurun_out/r2d_bench_2gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench_2gpu.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'bus_check', d['bus_check'], 'per rank', d['per_rank_ms_per_step'], 'e2e', (d.get('e2e') or {}).get('ms_per_step'))
print('cap', [(p.get('sources_per_gpu'), p.get('ms_per_block')) for p in (d['config'].get('capacity') or {}).get('passes', [])])"
timeout 600 $TR bench.py --gpus 2 --config 4 --steps 50 --warmup 5 > gpurun_out/r2d_bench_c4_2gpu.json 2> gpurun_out/r2d_bench_c4_2gpu.err; echo "2gpu c4 rc=$?"; tail -3 gpurun_out/r2d_bench_c4_2gpu.err; cut -c1-300 gpurun_out/r2d_bench_c4_2gpu.json
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench_c4_2gpu.json'))
print('c4 2gpu ms', d['ms_per_step'], d['bus_check'], d['per_rank_ms_per_step'], d['config']['shards_begin_end_sourcechannels'], d['kernel_ms_per_step_rank0'])"
timeout 300 python bench.py --config 4 --steps 30 --warmup 5 > gpurun_out/r2d_bench_c4_1gpu.json 2> gpurun_out/r2d_bench_c4_1gpu.err; echo "1gpu c4 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench_c4_1gpu.json'))
print('c4 1gpu ms', d['ms_per_step'], d['kernel_ms_per_step_rank0'])"
timeout 120 scratch/mb/issue_overlap_bench 2>&1 | tee gpurun_out/r2d_issue_overlap.txt
