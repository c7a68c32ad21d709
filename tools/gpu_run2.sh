#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_kernels.py -m gpu -q -s -k "micro_batch or preprocess" > gpurun_out/t6.log 2>&1; echo "t6 exit $?"; tail -n 6 gpurun_out/t6.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -n 5 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 --ops-csv gpurun_out/ops_resnet50.csv > gpurun_out/bench_r50.json 2> gpurun_out/bench_r50.err; echo "bench r50 exit $?"; tail -c 1500 gpurun_out/bench_r50.err; cat gpurun_out/bench_r50.json
timeout 900 python bench.py --model InceptionV3 --steps 10 --warmup 3 --no-cpu-baseline --ops-csv gpurun_out/ops_inception.csv > gpurun_out/bench_inc.json 2> gpurun_out/bench_inc.err; echo "bench inc exit $?"; tail -c 1500 gpurun_out/bench_inc.err; cat gpurun_out/bench_inc.json
