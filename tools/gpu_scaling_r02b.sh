#!/bin/bash
# Refresh of the 8-GPU figures with the final round-2 build (run as: gpurun --gpus 8 -- bash tools/gpu_scaling_r02b.sh):
# weak scaling of bench.py at N = 8 (ResNet50 batch 256 per GPU, InceptionV3 batch 128 per GPU attached) and strong scaling of
# InceptionV3 batch 128 image-sharded over 8 GPUs (BASELINE.json configs[3]).  The full 1 / 2 / 4 / 8 sweep is tools/gpu_scaling_r02.sh.
mkdir -p gpurun_out/scale_b
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29508 bench.py --gpus 8 --steps 20 --warmup 5 --sustain-seconds 1 --no-cpu-baseline \
    > gpurun_out/scale_b/weak_n8.json 2> gpurun_out/scale_b/weak_n8.err; echo "weak 8 rc $?"
$TR --nproc-per-node 8 --master-port 29608 bench.py --gpus 8 --model InceptionV3 --scaling strong --steps 20 --warmup 5 --sustain-seconds 0 \
    --no-cpu-baseline > gpurun_out/scale_b/strong_n8.json 2> gpurun_out/scale_b/strong_n8.err; echo "strong 8 rc $?"
tail -2 gpurun_out/scale_b/*.err | tail -8
