#!/bin/bash
# 8-GPU diagnosis: peer combine / NCCL / no combine, device-timed only
mkdir -p gpurun_out
for bus in ${BUSES:-peer nccl none}; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus 8 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e --bus $bus > gpurun_out/s8_$bus.json 2>gpurun_out/s8_$bus.err || tail -5 gpurun_out/s8_$bus.err
python - <<PY
import json
d=json.load(open("gpurun_out/s8_$bus.json"))
print("$bus", round(d["ms_per_step"],4), "host", d["host_enqueue_ms_per_step"], d["per_rank_ms_per_step"], {k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
