#!/bin/bash
# round-2 GPU pass G: A/B of the trace kernel without the separate alive flag against the previous commit's build.
mkdir -p gpurun_out
NEW=./metal-renderer_b200/bin/mrt_render; OLD=./metal-renderer_b200/prev/bin/mrt_render
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 64" "--scene many-lights --size 1920x1080 --seed 5 --frames 64" "--scene cornell-water --size 1920x1080 --seed 2 --frames 64"; do
  for rep in 1 2 3; do for B in $OLD $NEW; do echo "== $B $S"; MRT_WAVE_LOG=/dev/null timeout 600 $B $S --pipeline wavefront | tail -1; done; done
done
} 2>&1 | tee gpurun_out/g_ab.log
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_baseline_sizes.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/g_pytest.log
