#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --model InceptionV3 --steps 2 --warmup 3 --no-cpu-baseline --ncu-step"
$CMD > gpurun_out/ncu3_plain.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"pool|preprocess|stem" -c 8 -o gpurun_out/prof_pools $CMD > gpurun_out/ncu3.log 2>&1
echo "exit $?"
