#!/bin/bash
# config 4 at 1 / 2 / 4 GPUs of one box (the 1 -> 8 curve's lower points; 8 GPUs: r02n)
mkdir -p gpurun_out
python bench.py --config 4 --steps 50 --warmup 5 > gpurun_out/r02s_bench_c4_1gpu.json 2> gpurun_out/r02s_bench_c4_1gpu.err; echo "1 rc=$?"
for N in 2 4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --config 4 --steps 50 --warmup 5 > gpurun_out/r02s_bench_c4_${N}gpu.json 2> gpurun_out/r02s_bench_c4_${N}gpu.err; echo "$N rc=$?"
done
python - <<PY
import json
for N in (1, 2, 4):
    try:
        d = json.loads(open(f"gpurun_out/r02s_bench_c4_{N}gpu.json").read().strip().splitlines()[-1])
        print(N, "GPUs: total sources", d["config"]["total_sources"], "ms/block", round(d["ms_per_step"], 4), "real-time multiple", round(d["config"]["real_time_multiple"], 2), "value", round(d["value"]), d.get("bus_check"), d.get("per_rank_ms_per_step"))
    except Exception as e: print(N, "unreadable", e)
PY
