#!/bin/bash
# round-2 GPU pass C: A/B of the trace kernel with the loop-free leaf expansion (Bvh8Node::leafTris + shared spread table) and
# the merged alive ballot against the previous commit's build (metal-renderer_b200/prev), then the GPU test-suite on the new format.
mkdir -p gpurun_out
NEW=./metal-renderer_b200/bin/mrt_render; OLD=./metal-renderer_b200/prev/bin/mrt_render
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 64" "--scene cornell-water --size 1920x1080 --seed 2 --frames 64" \
         "--scene many-lights --size 1920x1080 --seed 5 --frames 64" "--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"; do
  for rep in 1 2; do for B in $OLD $NEW; do echo "== $B $S"; MRT_WAVE_LOG=/dev/null timeout 600 $B $S --pipeline wavefront | tail -1; done; done
  for B in $OLD $NEW; do echo "== graph $B $S"; timeout 600 $B $S --pipeline wavefront | tail -1; done
done
} 2>&1 | tee gpurun_out/c_ab.log
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/c_pytest.log
