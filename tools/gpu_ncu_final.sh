#!/bin/bash
# One ncu metrics pass over exactly one bench.py step (bracketed by cudaProfilerStart/Stop); run only after the
# same command exited 0 without ncu.
mkdir -p gpurun_out
MODEL=${1:-ResNet50}
CMD="python bench.py --model $MODEL --steps 2 --warmup 3 --no-cpu-baseline --no-second-model --sustain-seconds 0 --ncu-step"
$CMD > gpurun_out/ncu_plain_$MODEL.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain_$MODEL.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__t_bytes.sum \
    --clock-control none --csv --log-file gpurun_out/launches_$MODEL.csv $CMD > gpurun_out/ncu_$MODEL.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_$MODEL.log; wc -l gpurun_out/launches_$MODEL.csv
