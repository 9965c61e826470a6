#!/bin/bash
# round-1 GPU pass L: child interval clipped with min/max (one compare per child); SAH triangle-cost A/B
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_l.log
{ for i in 1 2; do timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront; done
  timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
  timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 --pipeline wavefront
  echo "cprim=0.5"; MRT_BVH8_CPRIM=0.5 timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  echo "cprim=0.15"; MRT_BVH8_CPRIM=0.15 timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
} 2>&1 | grep -v "^+" | tee gpurun_out/sweep_l.log
