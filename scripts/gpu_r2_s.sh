#!/bin/bash
# round-2 GPU pass S (2 GPUs): where the C driver's wall time goes with --reduce peer (rank_setup_ms / rank_teardown_ms)
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128"
{ for k in 1 2; do for red in nccl peer; do timeout 300 $R $S --devices 0,1 --shard samples --reduce $red | tail -1 | cut -c150-; done; done; } 2>&1 | tee gpurun_out/s_phases.log
