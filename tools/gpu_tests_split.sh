#!/bin/bash
# Runs the GPU tests in separate processes (a faulting kernel poisons its CUDA
# context) and collects logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?"; tail -n 25 gpurun_out/$name.log; }
run t1_kernels_basic python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "preprocess or cuda_core"
run t2_kernels_tc python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tcgen05"
run t3_parity_basic python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "fp32_check or cuda_core or folding"
run t4_parity_tc python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "tcgen05 or micro_batch"
run t5_models_api python -m pytest tests/test_gpu_models_api.py -m gpu -q -s
