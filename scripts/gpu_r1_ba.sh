#!/bin/bash
# round-1 GPU pass BA: ncu of one steady-state trace launch on the 10M-triangle 4K workload with the selected (64-bin) tree
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-10m --size 3840x2160 --seed 4 --frames 4 --pipeline wavefront"
timeout 300 $R $S > gpurun_out/plain_ba.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 20 -c 1 -f -o gpurun_out/prof_ba_trace_10m $R $S > gpurun_out/ncu_ba.log 2>&1
ls -la gpurun_out/prof_ba_trace_10m.ncu-rep
