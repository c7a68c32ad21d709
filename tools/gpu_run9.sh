#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "all gpu tests exit $?"; tail -n 6 gpurun_out/t_all.log
for M in ResNet50 InceptionV3; do
timeout 900 python bench.py --model $M --steps 10 --warmup 3 --no-cpu-baseline --ops-csv gpurun_out/ops_$M.csv > gpurun_out/bench_$M.json 2> gpurun_out/bench_$M.err; echo "bench exit $?"; tail -c 800 gpurun_out/bench_$M.err; python -c "
import json; d=json.load(open('gpurun_out/bench_$M.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']); print({k:(round(v['ms'],3), round(v['tflops'],1), round(v['gbs'])) for k,v in d['kernels'].items()})"
done
