#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "stem" > gpurun_out/t11.log 2>&1; echo "t11 exit $?"; tail -n 8 gpurun_out/t11.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "tcgen05" > gpurun_out/t12.log 2>&1; echo "t12 exit $?"; grep -E "rel|passed|failed|Error" gpurun_out/t12.log | tail -n 12
for M in ResNet50 InceptionV3; do
timeout 900 python bench.py --model $M --steps 10 --warmup 3 --no-cpu-baseline --ops-csv gpurun_out/ops_$M.csv > gpurun_out/bench_$M.json 2> gpurun_out/bench_$M.err; echo "bench exit $?"; tail -c 800 gpurun_out/bench_$M.err; python -c "
import json; d=json.load(open('gpurun_out/bench_$M.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']); print({k:(round(v['ms'],3), round(v['tflops'],1), round(v['gbs'])) for k,v in d['kernels'].items()})"
done
