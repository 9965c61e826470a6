#!/bin/bash
# round-1 GPU pass W8: 8-GPU and 4-GPU bench under torchrun (as the driver launches it)
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for n in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
  bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_r1_w_${n}gpu.json 2> gpurun_out/bench_r1_w_${n}gpu.err
tail -c 600 gpurun_out/bench_r1_w_${n}gpu.json; tail -3 gpurun_out/bench_r1_w_${n}gpu.err
done
