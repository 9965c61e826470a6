#!/bin/bash
# round-2 GPU pass F (8 GPUs): bench.py under torchrun at N = 8 on config 3 (weak 128 spp per GPU + strong 128 spp total + parity of
# the all-reduced image) and on config 4 (10 M triangles, 3840x2160, 8 x 32 = 256 spp per step), and the C driver on 8 GPUs.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | tee gpurun_out/f8_gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29521 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/f8_bench_8gpu.json 2> gpurun_out/f8_bench_8gpu.err; tail -c 2500 gpurun_out/f8_bench_8gpu.json; tail -3 gpurun_out/f8_bench_8gpu.err
timeout 1200 $TR --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 3 --workload spheres-10m-4k --frames 32 > gpurun_out/f8_bench_8gpu_cfg4.json 2> gpurun_out/f8_bench_8gpu_cfg4.err; tail -c 2500 gpurun_out/f8_bench_8gpu_cfg4.json; tail -3 gpurun_out/f8_bench_8gpu_cfg4.err
R=./metal-renderer_b200/bin/mrt_render
{
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128 --depth 8"
timeout 300 $R $S --out gpurun_out/f8_1gpu.pfm
timeout 300 $R $S --devices 0,1,2,3,4,5,6,7 --shard samples --out gpurun_out/f8_8gpu_samples.pfm
timeout 300 $R $S --devices 0,1,2,3,4,5,6,7 --shard rows --out gpurun_out/f8_8gpu_rows.pfm
python scripts/compare_pfm.py gpurun_out/f8_8gpu_rows.pfm gpurun_out/f8_1gpu.pfm
python scripts/compare_pfm.py gpurun_out/f8_8gpu_samples.pfm gpurun_out/f8_1gpu.pfm
rm -f gpurun_out/f8_*.pfm
} 2>&1 | tee gpurun_out/f8_c_driver_8gpu.log
