#!/bin/bash
# 2 GPUs: peer tests on one device, then the default bench (re-export reuses the exchange block; teardown disconnects, barriers, frees)
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_peer_reduce.py tests/test_gpu_sharded.py -m gpu -q 2>&1 | tail -3 ) | tee gpurun_out/za_tests.log
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/za_bench_2gpu.json 2> gpurun_out/za_bench_2gpu.err; echo "rc=$?"; tail -4 gpurun_out/za_bench_2gpu.err; tail -c 200 gpurun_out/za_bench_2gpu.json
for red in nccl peer; do ./metal-renderer_b200/bin/mrt_render --scene spheres-1m --size 1920x1080 --seed 3 --frames 32 --devices 0,1 --reduce $red | tail -1 | cut -c150-420; done
