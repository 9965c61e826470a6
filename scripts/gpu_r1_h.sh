#!/bin/bash
# round-1 GPU pass H (dynamic work items): wide BVH with PRMT plane decode + single-triangle steps — parity, sweep, ncu
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu_h.log
S="--scene spheres-1m --size 1920x1080 --frames 16 --seed 3"
for mb in 8; do for lm in 8; do for rf in 24; do
  echo "wide minb=$mb leafmin=$lm refill=$rf"; MRT_TRACE_MINBLOCKS=$mb MRT_TRACE_LEAFMIN=$lm MRT_TRACE_REFILL=$rf timeout 300 $R $S --pipeline wavefront
done; done; done 2>&1 | tee gpurun_out/sweep_h.log
echo "binary"; MRT_BVH_WIDTH=2 timeout 300 $R $S --pipeline wavefront 2>&1 | tee -a gpurun_out/sweep_h.log
timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront 2>&1 | tee -a gpurun_out/sweep_h.log
timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 --pipeline wavefront 2>&1 | tee -a gpurun_out/sweep_h.log
timeout 600 $R --scene spheres-10m --size 3840x2160 --frames 4 --seed 4 --pipeline wavefront 2>&1 | tee -a gpurun_out/sweep_h.log
timeout 300 $R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 --pipeline wavefront > gpurun_out/plain_wf2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 2 -c 1 -o gpurun_out/prof_wf_h \
    $R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 --pipeline wavefront > gpurun_out/ncu_wf_trace.log 2>&1
ls -la gpurun_out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_r1_h.json 2> gpurun_out/bench_r1_h.err; tail -c 2500 gpurun_out/bench_r1_h.json; tail -3 gpurun_out/bench_r1_h.err
