#!/bin/bash
# ncu evidence for bench.py's step: launch list (all kernels) + one full capture of the conv kernel.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 470 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc -s 40 -c 3 -o gpurun_out/prof_conv_tc $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"
tail -n 3 gpurun_out/ncu_list.log gpurun_out/ncu_full.log
