#!/bin/bash
# final round-2 session on one GPU: tests, smoke, contract line, reference arm, other configs, ncu launch list and full captures
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02z_gpu_tests.txt 2>&1; tail -n 3 gpurun_out/r02z_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02z_smoke.txt 2>&1; tail -n 3 gpurun_out/r02z_smoke.txt
python bench.py > gpurun_out/r02z_bench_default.json 2> gpurun_out/r02z_bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02z_bench_reference.json 2> gpurun_out/r02z_bench_reference.err; echo "ref rc=$?"
python bench.py --mode throughput --no-cpu-baseline --no-e2e > gpurun_out/r02z_bench_throughput.json 2>/dev/null
for c in 1 2 4 5; do python bench.py --config $c > gpurun_out/r02z_bench_config$c.json 2> gpurun_out/r02z_bench_config$c.err; echo "config $c rc=$?"; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02z_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02z_ncu_bench.log 2>&1
python scratch/one_render.py 8 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render -s 5 -c 1 -f -o gpurun_out/r02z_k_render python scratch/one_render.py 8 > gpurun_out/r02z_ncu_render.log 2>&1
JPLG_FUSED_PUSH=0 ncu --set full --clock-control none --import-source on -k regex:k_render -s 5 -c 1 -f -o gpurun_out/r02z_k_render_nopush python scratch/one_render.py 8 > gpurun_out/r02z_ncu_render2.log 2>&1
tail -n 1 gpurun_out/r02z_ncu_render.log gpurun_out/r02z_ncu_render2.log
