#!/bin/bash
# round-1 GPU pass M: vote ratio (triangle step when 16*nT >= r*nN) and L1 prefetch of the first leaf triangle
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for pf in 0 1; do for tr in 0 5 6 8; do
  echo "prefetch=$pf triratio=$tr"
  MRT_TRACE_PREFETCH=$pf MRT_TRACE_TRIRATIO=$tr timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_TRACE_PREFETCH=$pf MRT_TRACE_TRIRATIO=$tr timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
done; done 2>&1 | grep -v "^+" | tee gpurun_out/sweep_m.log
for lm in 6 10; do echo "leafmin=$lm"; MRT_TRACE_LEAFMIN=$lm timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront; done 2>&1 | grep -v "^+" | tee -a gpurun_out/sweep_m.log
MRT_TRACE_PREFETCH=1 MRT_TRACE_TRIRATIO=6 timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_m.log
