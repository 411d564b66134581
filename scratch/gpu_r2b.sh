#!/bin/bash
# r02b: parity tests (fused push role, new config-shape tests), A/B of the push / window variants, the new bench.py paths
mkdir -p gpurun_out
# fail fast: a hung kernel must not burn the lease
timeout 240 python __graft_entry__.py --smoke > gpurun_out/r2b_smoke.txt 2>&1; rc=$?; tail -4 gpurun_out/r2b_smoke.txt
if [ $rc -ne 0 ]; then echo "SMOKE FAILED rc=$rc"; exit 1; fi
timeout 900 python -m pytest tests -m gpu -x -q -p no:hypothesispytest --timeout=240 2>&1 | tail -15 | tee gpurun_out/r2b_tests.txt
if ! grep -q " passed" gpurun_out/r2b_tests.txt; then echo "TESTS DID NOT FINISH"; exit 1; fi
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], round(d["ms_per_step"], 5), d.get("bus_crc32"), round(d["roofline"]["avg_launch_ms"] * 1e3, 1), {k: round(x * 1e3, 1) for k, x in d["roofline"]["kernel_ms_per_step"].items()}, "frac", round(d["roofline"]["frac"], 3))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
Q="--steps 300 --warmup 20 --no-cpu-baseline --no-e2e --no-capacity --no-offline"
for rep in 1; do
  for sched in realtime packed; do
    timeout 200 python bench.py $Q --schedule $sched > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "new/$sched" gpurun_out/ab.json
    JPLG_FUSED_PUSH=0 timeout 200 python bench.py $Q --schedule $sched > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "nofuse-bulkpush/$sched" gpurun_out/ab.json
    JPLG_FUSED_PUSH=0 JPLG_PUSH_VEC4=1 timeout 200 python bench.py $Q --schedule $sched > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "nofuse-vec4push/$sched" gpurun_out/ab.json
    JPLG_LIB_PATH=$PWD/scratch/libjplg_wintma.so timeout 200 python bench.py $Q --schedule $sched > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "wintma/$sched" gpurun_out/ab.json
  done
done 2>&1 | tee gpurun_out/r2b_ab.txt
timeout 200 python bench.py $Q --mode throughput > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; show "throughput/realtime" gpurun_out/ab.json | tee -a gpurun_out/r2b_ab.txt
# the full default line (capacity pass, offline schedule, e2e, CPU baseline)
timeout 900 python bench.py > gpurun_out/r2b_bench_default.json 2> gpurun_out/r2b_bench_default.err; echo "default bench rc=$?"; tail -5 gpurun_out/r2b_bench_default.err
python -c "
import json; d=json.load(open('gpurun_out/r2b_bench_default.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', (d.get('e2e') or {}).get('ms_per_step'), 'cap', json.dumps(d['config'].get('capacity'))[:1500])
print('offline', d['config'].get('offline_schedule')); print('roofline', {k:v for k,v in d['roofline'].items() if k in ('bound','achieved','peak','frac','avg_launch_ms')}); print('cpu', d.get('cpu_baseline'))"
for c in 1 2 4 5; do timeout 300 python bench.py --config $c --steps 50 --warmup 5 > gpurun_out/r2b_bench_c$c.json 2> gpurun_out/r2b_bench_c$c.err; echo "config $c rc=$?"; tail -2 gpurun_out/r2b_bench_c$c.err; cut -c1-600 gpurun_out/r2b_bench_c$c.json; done
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r2b_bench_ref.json 2> gpurun_out/r2b_bench_ref.err; echo "ref rc=$?"; cut -c1-1200 gpurun_out/r2b_bench_ref.json
JPLG_LIB_PATH=$PWD/scratch/libjplg_prof.so timeout 200 python scratch/render_prof.py 2>&1 | tail -8
