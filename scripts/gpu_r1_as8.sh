#!/bin/bash
# round-1 GPU pass AS8: 8-GPU bench with the selected trees — config 3 (default) and config 4 (10 M triangles, 4K; 8 GPUs x 32
# frames per step = the configuration's 256 spp), under torchrun as the driver launches it
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29628 \
  bench.py --gpus 8 --workload spheres-10m-4k --frames 32 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_as_8gpu_cfg4.json 2> gpurun_out/bench_r1_as_8gpu_cfg4.err
tail -c 700 gpurun_out/bench_r1_as_8gpu_cfg4.json; tail -3 gpurun_out/bench_r1_as_8gpu_cfg4.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29618 \
  bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_as_8gpu.json 2> gpurun_out/bench_r1_as_8gpu.err
tail -c 700 gpurun_out/bench_r1_as_8gpu.json; tail -3 gpurun_out/bench_r1_as_8gpu.err
