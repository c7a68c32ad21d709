#!/bin/bash
# Round-2 multi-GPU evidence on ONE 8-GPU box (run as: gpurun --gpus 8 -- bash tools/gpu_scaling_r02.sh):
#   weak scaling of bench.py at N = 2, 4, 8 (ResNet50 batch 256 per GPU, InceptionV3 batch 128 per GPU attached),
#   strong scaling of InceptionV3 batch 128 image-sharded over N = 1, 2, 4, 8 (BASELINE.json configs[3]),
#   the job mix of configs[4] on 8 worker processes.
mkdir -p gpurun_out/scale
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29500 + n)) bench.py --gpus $n --steps 20 --warmup 5 --sustain-seconds 1 --no-cpu-baseline \
      > gpurun_out/scale/weak_n$n.json 2> gpurun_out/scale/weak_n$n.err; echo "weak $n rc $?"
done
python bench.py --model InceptionV3 --scaling strong --steps 20 --warmup 5 --sustain-seconds 1 --no-cpu-baseline > gpurun_out/scale/strong_n1.json 2> gpurun_out/scale/strong_n1.err; echo "strong 1 rc $?"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600 + n)) bench.py --gpus $n --model InceptionV3 --scaling strong --steps 20 --warmup 5 --sustain-seconds 1 \
      --no-cpu-baseline > gpurun_out/scale/strong_n$n.json 2> gpurun_out/scale/strong_n$n.err; echo "strong $n rc $?"
done
timeout 600 python tools/job_mix.py --gpus 8 --jobs-per-model 4 --out gpurun_out/scale/job_mix_n8.json > /dev/null 2> gpurun_out/scale/job_mix_n8.err; echo "job mix rc $?"
tail -2 gpurun_out/scale/*.err | tail -20
