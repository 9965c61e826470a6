#!/bin/bash
# round-1 GPU pass P: full GPU suite (display + checkpoint included), shade kernel at 3 vs 4 blocks per SM
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_p.log
for mb in 3 4 3 4; do
  echo "shade_minblocks=$mb"
  MRT_SHADE_MINBLOCKS=$mb timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_SHADE_MINBLOCKS=$mb timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
done 2>&1 | grep -v "^+" | tee gpurun_out/sweep_p.log
MRT_SHADE_MINBLOCKS=4 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "production or config1 or nans" 2>&1 | tail -3
