#!/bin/bash
# N-GPU session: bare H2D probe at 1 / 2 / 4 / N ranks, config 3 contract line at N, config 4 at N.   usage: gpu_r2n.sh N
N=${1:-8}
mkdir -p gpurun_out
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "${@:2}"; }
for r in 1 2 4 8; do
  if [ $r -le $N ]; then timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $r --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) scratch/h2d_probe_multi.py > gpurun_out/r02n_h2d_probe_${r}.json 2> gpurun_out/r02n_h2d_probe_${r}.err; tail -n 1 gpurun_out/r02n_h2d_probe_${r}.json; fi
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r02n_bench_c3_${N}gpu.json 2> gpurun_out/r02n_bench_c3_${N}gpu.err; echo "c3 rc=$?"; tail -n 2 gpurun_out/r02n_bench_c3_${N}gpu.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --config 4 --steps 50 --warmup 5 > gpurun_out/r02n_bench_c4_${N}gpu.json 2> gpurun_out/r02n_bench_c4_${N}gpu.err; echo "c4 rc=$?"; tail -n 2 gpurun_out/r02n_bench_c4_${N}gpu.err
python - <<PY
import json
for f in ("gpurun_out/r02n_bench_c3_${N}gpu.json", "gpurun_out/r02n_bench_c4_${N}gpu.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "value", d["value"], "ms", round(d["ms_per_step"], 4), "bus_check", d.get("bus_check"), "e2e", (d.get("e2e") or {}).get("ms_per_step"), (d.get("e2e") or {}).get("value"))
    except Exception as e: print(f, "unreadable", e)
PY
