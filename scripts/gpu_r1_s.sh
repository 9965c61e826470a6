#!/bin/bash
# round-1 GPU pass S: postponed triangles in the trace kernel (MRT_TRACE_DEFER) — parity and threshold sweep
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -k "tuning" 2>&1 | tail -3
MRT_TRACE_DEFER=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x -k "not tuning" 2>&1 | tail -3
{ echo "defer=0"; MRT_TRACE_COUNT=1 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 16 --pipeline wavefront
  $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
for th in 8 12 16 20 24 33; do for lm in 8 16; do
  echo "defer=1 thresh=$th leafmin=$lm"
  export MRT_TRACE_DEFER=1 MRT_TRACE_TRITHRESH=$th MRT_TRACE_LEAFMIN=$lm
  MRT_TRACE_COUNT=1 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 16 --pipeline wavefront
  $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
done; done; } 2>&1 | grep -v "^+" | tee gpurun_out/sweep_s.log
