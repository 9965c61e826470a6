#!/bin/bash
# round-1 GPU pass K: slot groups on separate streams (trace tails overlapped by the other group's kernels)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_k.log
for sp in 1 2 3 4; do
  echo "split=$sp"
  MRT_WF_SPLIT=$sp timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_WF_SPLIT=$sp timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_WF_SPLIT=$sp timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
  MRT_WF_SPLIT=$sp timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 --pipeline wavefront
done 2>&1 | grep -v "^+" | tee gpurun_out/sweep_k.log
for mb in 6 7; do echo "split=2 minb=$mb"; MRT_TRACE_MINBLOCKS=$mb MRT_WF_SPLIT=2 timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront; done 2>&1 | grep -v "^+" | tee -a gpurun_out/sweep_k.log
