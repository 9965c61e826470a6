#!/bin/bash
# round-1 GPU pass Z: ncu of a steady-state k_wf_shade launch in its final form (64 registers, 4 blocks per SM)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
timeout 300 $R $S > gpurun_out/plain_z.log 2>&1 && cat gpurun_out/plain_z.log && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 31 -c 1 -o gpurun_out/prof_wf_z_shade $R $S > gpurun_out/ncu_z_shade.log 2>&1
ls -la gpurun_out | tail -4
