#!/bin/bash
# ncu evidence for the final round-1 build: every launch of ONE bench step (time, DRAM bytes, L2 bytes, tensor pipe),
# and one full capture of the dominant kernel.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --ncu-step"
$CMD > gpurun_out/ncu2_plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__t_bytes.sum --clock-control none --csv --log-file gpurun_out/launches2.csv $CMD > gpurun_out/ncu2_list.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/ncu2_plain2.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:conv_tc2 -s 6 -c 4 -o gpurun_out/prof_conv_final $CMD > gpurun_out/ncu2_full.log 2>&1
echo "full capture exit $?"
