#!/bin/bash
# r02a: parity tests on the 5-role k_render, then same-box A/B of library builds, role profile, FP32x2 microbenchmark
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 1500 python -m pytest tests -m gpu -x -q -p no:hypothesispytest 2>&1 | tail -15 | tee gpurun_out/r2a_tests.txt
for rep in 1 2; do
for v in old new mixfma; do
if [ $v = new ]; then unset JPLG_LIB_PATH; else export JPLG_LIB_PATH=$PWD/scratch/libjplg_$v.so; fi
for extra in "" "--no-push-ahead --no-pan-overlap"; do
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e $extra > gpurun_out/ab_$v.json 2>gpurun_out/ab_$v.err || tail -3 gpurun_out/ab_$v.err
python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v.json"))
print("$v", "$extra", round(d["ms_per_step"],5), d["bus_crc32"], round(d["roofline"]["avg_launch_ms"]*1e3,1), {k:round(x*1e3,1) for k,x in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
done
done 2>&1 | tee gpurun_out/r2a_ab.txt
unset JPLG_LIB_PATH
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e --mode throughput > gpurun_out/ab_thr.json 2>gpurun_out/ab_thr.err || tail -3 gpurun_out/ab_thr.err
python -c "
import json
d=json.load(open('gpurun_out/ab_thr.json'))
print('throughput', round(d['ms_per_step'],5), round(d['roofline']['avg_launch_ms']*1e3,1), {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms_per_step'].items()})" | tee -a gpurun_out/r2a_ab.txt
JPLG_LIB_PATH=$PWD/scratch/libjplg_prof.so timeout 200 python scratch/render_prof.py 2>&1 | tail -8 | tee gpurun_out/r2a_prof.txt
timeout 120 scratch/mb/fp32x2_bench 2>&1 | tee gpurun_out/r2a_fp32x2.txt | tail -45
