#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --model InceptionV3 --steps 2 --warmup 3 --no-cpu-baseline --no-second-model --sustain-seconds 0 --ncu-step"
$CMD > gpurun_out/ncu_pool_plain.log 2>&1 || { echo "plain failed"; exit 1; }
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"avgpool3|maxpool3" -c 6 -o gpurun_out/prof_pools_r01 $CMD > gpurun_out/ncu_pool.log 2>&1
echo "exit $?"
