#!/bin/bash
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
for v in 24; do
MRT_TRACE_TQ=$v timeout 300 $R $S > /dev/null 2>&1 && \
MRT_TRACE_TQ=$v timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -f -o gpurun_out/prof_aj_tq$v $R $S > gpurun_out/ncu_aj_tq$v.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
