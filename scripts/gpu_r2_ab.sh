#!/bin/bash
# A/B of the working tree's build against metal-renderer_b200/prev (the build before the change), alternating, graph schedule:
# device ms of 64-frame batches on the four BASELINE scenes, then the parity tests that cover the traversal.
mkdir -p gpurun_out
NEW=./metal-renderer_b200/bin/mrt_render; OLD=./metal-renderer_b200/prev/bin/mrt_render
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 64" "--scene many-lights --size 1920x1080 --seed 5 --frames 64" \
         "--scene cornell-water --size 1920x1080 --seed 2 --frames 64" "--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"; do
  for rep in 1 2 3; do for B in $OLD $NEW; do echo -n "$B $S: "; timeout 600 $B $S --pipeline wavefront | tail -1 | grep -o '"device_ms": [0-9.]*, "kernel_ms": [0-9.]*'; done; done
done
} 2>&1 | tee gpurun_out/ab_${1:-x}.log
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_baseline_sizes.py -m gpu -x -q 2>&1 | tail -3 ) | tee gpurun_out/ab_${1:-x}_pytest.log
