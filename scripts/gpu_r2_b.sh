#!/bin/bash
# round-2 GPU pass B (2 GPUs): the C driver with NCCL (--devices 0,1: rows mode must be bit-identical to one GPU, samples mode
# within rRMSE 1e-5), bench.py under torchrun at N = 2 (weak + strong + parity of the all-reduced image), media scene parity.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | tee gpurun_out/b_gpus.txt
R=./metal-renderer_b200/bin/mrt_render
{
for S in "--scene cornell-water --size 640x360 --seed 2 --frames 16 --depth 8" "--scene spheres-1m --size 1920x1080 --seed 3 --frames 16 --depth 8"; do
  tag=$(echo $S | awk '{print $2}')
  echo "== $S"
  timeout 300 $R $S --out gpurun_out/b_${tag}_1gpu.pfm
  timeout 300 $R $S --devices 0,1 --shard rows --out gpurun_out/b_${tag}_2gpu_rows.pfm
  timeout 300 $R $S --devices 0,1 --shard samples --out gpurun_out/b_${tag}_2gpu_samples.pfm
  python scripts/compare_pfm.py gpurun_out/b_${tag}_2gpu_rows.pfm gpurun_out/b_${tag}_1gpu.pfm
  python scripts/compare_pfm.py gpurun_out/b_${tag}_2gpu_samples.pfm gpurun_out/b_${tag}_1gpu.pfm
  cmp gpurun_out/b_${tag}_2gpu_rows.pfm gpurun_out/b_${tag}_1gpu.pfm && echo "rows mode: PFM bit-identical to the 1-GPU render"
done
rm -f gpurun_out/b_*.pfm
} 2>&1 | tee gpurun_out/b_c_driver_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 --frames 32 \
   > gpurun_out/b_bench_2gpu.json 2> gpurun_out/b_bench_2gpu.err; tail -c 3000 gpurun_out/b_bench_2gpu.json; tail -5 gpurun_out/b_bench_2gpu.err
( CUDA_VISIBLE_DEVICES=0 timeout 600 python -m pytest tests/test_gpu_baseline_sizes.py -k media -x -q 2>&1 | tail -5 ) | tee gpurun_out/b_pytest_media.log
RX=./metal-renderer_b200/bin/mrt_render_exp
S="--scene spheres-10m --size 3840x2160 --seed 4 --frames 8 --pipeline wavefront"
{ for v in 0 3; do echo "== 10M MRT_WF_SORT=$v (trace time from the wave log)"; CUDA_VISIBLE_DEVICES=1 MRT_WF_SORT=$v MRT_WAVE_LOG=gpurun_out/b_waves_10m_sort$v.csv timeout 600 $RX $S | tail -1; done; } 2>&1 | tee gpurun_out/b_10m_sort.log
