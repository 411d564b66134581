#!/bin/bash
mkdir -p gpurun_out
cp jplspatial_b200/libjplg.so /tmp/prod.so
cp scratch/libjplg_prof.so jplspatial_b200/libjplg.so
python scratch/render_prof.py > gpurun_out/r2t_prof.txt 2>&1
JPLG_FUSED_PUSH=0 python scratch/render_prof.py > gpurun_out/r2t_prof_nofuse.txt 2>&1
cp /tmp/prod.so jplspatial_b200/libjplg.so
cat gpurun_out/r2t_prof.txt; echo "--- separate push kernel:"; cat gpurun_out/r2t_prof_nofuse.txt
