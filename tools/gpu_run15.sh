#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "all gpu tests exit $?"; tail -n 6 gpurun_out/t_all.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; tail -c 600 gpurun_out/bench_full.err; python -c "
import json; d=json.load(open('gpurun_out/bench_full.json')); print({k:d[k] for k in ('value','ms_per_step','e2e','gpu_launches','clocks','cpu_baseline')}); print(d['roofline'])"
