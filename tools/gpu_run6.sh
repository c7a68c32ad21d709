#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tcgen05" > gpurun_out/t9.log 2>&1; echo "t9 exit $?"; tail -n 8 gpurun_out/t9.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "tcgen05 or micro_batch" > gpurun_out/t10.log 2>&1; echo "t10 exit $?"; grep -E "rel|passed|failed|Error" gpurun_out/t10.log | tail -n 12
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --ops-csv gpurun_out/ops_resnet50_v3.csv > gpurun_out/bench_r50_v3.json 2> gpurun_out/bench_r50_v3.err; echo "bench exit $?"; tail -c 800 gpurun_out/bench_r50_v3.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r50_v3.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']); print({k:(round(v['ms'],3), round(v['tflops'],1), round(v['gbs'])) for k,v in d['kernels'].items()})"
timeout 900 python bench.py --model InceptionV3 --steps 10 --warmup 3 --no-cpu-baseline --ops-csv gpurun_out/ops_inception_v3.csv > gpurun_out/bench_inc_v3.json 2> gpurun_out/bench_inc_v3.err; echo "bench exit $?"; tail -c 800 gpurun_out/bench_inc_v3.err; python -c "
import json; d=json.load(open('gpurun_out/bench_inc_v3.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']); print({k:(round(v['ms'],3), round(v['tflops'],1), round(v['gbs'])) for k,v in d['kernels'].items()})"
