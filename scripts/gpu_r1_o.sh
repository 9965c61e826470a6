#!/bin/bash
# round-1 GPU pass O: display + checkpoint parity, full suite, ncu of a steady-state wave (trace and shade kernels)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_o.log
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
timeout 300 $R $S --out gpurun_out/spheres_o.pfm --tone --srgb --checkpoint gpurun_out/spheres_o.ckpt > gpurun_out/plain_o.log 2>&1 && cat gpurun_out/plain_o.log && \
timeout 300 $R $S --resume gpurun_out/spheres_o.ckpt | tee -a gpurun_out/plain_o.log && rm -f gpurun_out/spheres_o.ckpt gpurun_out/spheres_o.pfm && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -o gpurun_out/prof_wf_o_trace $R $S > gpurun_out/ncu_o_trace.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 31 -c 1 -o gpurun_out/prof_wf_o_shade $R $S > gpurun_out/ncu_o_shade.log 2>&1
ls -la gpurun_out | tail -8
