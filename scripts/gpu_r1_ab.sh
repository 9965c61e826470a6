#!/bin/bash
# round-1 GPU pass AB: megakernel on the wide traversal core (like-for-like with the wavefront pipeline) + A/B of the
# sign-bit clamp in the node test (lib_ab/old.so = default build, lib_ab/new.so = -DMRT_SIGN_CLAMP)
set -x
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8 ) 2>&1 | tee gpurun_out/pytest_gpu_ab.log
R=./metal-renderer_b200/bin/mrt_render
for sc in spheres-1m cornell-water; do
for v in "MRT_FUSED_BVH=2 --pipeline fused" "MRT_FUSED_BVH=8 --pipeline fused" "MRT_FUSED_BVH=8 MRT_FUSED_MINBLOCKS=3 --pipeline fused" "MRT_X=0 --pipeline wavefront"; do
  echo "== $sc $v"
  env ${v% --pipeline*} timeout 600 $R --scene $sc --size 1920x1080 --seed 3 --frames 16 --pipeline ${v##*--pipeline }
done; done 2>&1 | grep -v "^+" | tee gpurun_out/megakernel_ab.log
bash scripts/ab_shade.sh 2>&1 | grep -v "^+" | tee gpurun_out/ab_signclamp.log
