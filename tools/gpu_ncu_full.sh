#!/bin/bash
# One `ncu --set full` capture of the dominant kernels of a bench.py step (after the same command exited 0 without ncu).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-second-model --sustain-seconds 0 --ncu-step"
$CMD > gpurun_out/ncu_full_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
# launches 1-8 of the profiled step's conv kernels: stem (v6), conv2 block 1 (1x1 shortcut, 1x1 reduce, 3x3 v6, 1x1 expand + residual), ...
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:conv_tc -c 12 -o gpurun_out/prof_conv_r01_final $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; tail -2 gpurun_out/ncu_full.log | cut -c1-200; ls -la gpurun_out/prof_conv_r01_final.ncu-rep
