#!/bin/bash
# round-1 GPU pass N2: 2-GPU bench under torchrun (NCCL all-reduce of the sum images), then 1-GPU for comparison
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_n2_2gpu.json 2> gpurun_out/bench_r1_n2_2gpu.err
tail -c 1500 gpurun_out/bench_r1_n2_2gpu.json; tail -5 gpurun_out/bench_r1_n2_2gpu.err
timeout 900 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_n2_1gpu.json 2> gpurun_out/bench_r1_n2_1gpu.err
tail -c 1500 gpurun_out/bench_r1_n2_1gpu.json; tail -5 gpurun_out/bench_r1_n2_1gpu.err
