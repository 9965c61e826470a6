#!/bin/bash
# round-2 GPU pass T (N GPUs): C driver, samples sharding, NCCL vs peer kernel with the peer pages touched before the timed region
N=${1:-8}; DEV=$(seq -s, 0 $((N-1)))
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128"
{ for k in 1 2 3; do for red in nccl peer; do timeout 300 $R $S --devices $DEV --shard samples --reduce $red --out gpurun_out/t_$red.pfm | tail -1 | cut -c150-; done; done
python scripts/compare_pfm.py gpurun_out/t_peer.pfm gpurun_out/t_nccl.pfm; } 2>&1 | tee gpurun_out/t_c_driver_${N}gpu.log
rm -f gpurun_out/t_*.pfm
