#!/bin/bash
# 2 GPUs: config 3 with bus_check, config 4 with bus_check
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/r2g_bench_2gpu.json 2> gpurun_out/r2g_bench_2gpu.err; echo "2gpu c3 rc=$?"; tail -3 gpurun_out/r2g_bench_2gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench_2gpu.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'bus_check', d['bus_check'], 'per rank', d['per_rank_ms_per_step'], 'e2e', (d.get('e2e') or {}).get('ms_per_step'))
print('cap', [(p.get('sources_per_gpu'), p.get('ms_per_block')) for p in (d['config'].get('capacity') or {}).get('passes', [])])"
timeout 600 $TR bench.py --gpus 2 --config 4 --steps 50 --warmup 5 > gpurun_out/r2g_bench_c4_2gpu.json 2> gpurun_out/r2g_bench_c4_2gpu.err; echo "2gpu c4 rc=$?"; tail -3 gpurun_out/r2g_bench_c4_2gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench_c4_2gpu.json'))
print('c4 2gpu ms', d['ms_per_step'], d['bus_check'], d['per_rank_ms_per_step'], d['config']['shards_begin_end_sourcechannels'], d['kernel_ms_per_step_rank0'])"
