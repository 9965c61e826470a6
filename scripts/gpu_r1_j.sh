#!/bin/bash
# round-1 GPU pass J: work-claim granularity (tile per warp / 256-pixel region per block / one global counter)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_j.log
for rs in 32 256 0; do
  echo "region=$rs"
  MRT_WF_REGION=$rs MRT_WAVE_LOG=gpurun_out/waves_j_spheres_$rs.csv timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_WF_REGION=$rs timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront
  MRT_WF_REGION=$rs MRT_WAVE_LOG=gpurun_out/waves_j_water_$rs.csv timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
  MRT_WF_REGION=$rs timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
  MRT_WF_REGION=$rs timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 --pipeline wavefront
done 2>&1 | grep -v "^+" | tee gpurun_out/sweep_j.log
MRT_WF_REGION=256 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "production or config1" 2>&1 | tail -3
MRT_WF_REGION=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "production or config1" 2>&1 | tail -3
