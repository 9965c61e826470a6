#!/bin/bash
# round-1 GPU pass AB2: megakernel on the binary tree / on the wide traversal core / wavefront, same scenes
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for sc in spheres-1m cornell-water; do
for v in "MRT_FUSED_BVH=2 --pipeline fused" "MRT_FUSED_BVH=8 --pipeline fused" "MRT_FUSED_BVH=8 MRT_FUSED_MINBLOCKS=3 --pipeline fused" "MRT_X=0 --pipeline wavefront"; do
  echo "== $sc $v"
  env ${v% --pipeline*} timeout 600 $R --scene $sc --size 1920x1080 --seed 3 --frames 16 --pipeline ${v##*--pipeline }
  env ${v% --pipeline*} timeout 600 $R --scene $sc --size 1920x1080 --seed 3 --frames 16 --pipeline ${v##*--pipeline }
done; done 2>&1 | grep -v "^+" | tee gpurun_out/megakernel_ab.log
