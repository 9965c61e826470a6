#!/bin/bash
# round-1 GPU pass AC: L2 access-policy window over the wide tree + triangle records (MRT_L2_PERSIST), A/B by environment
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 0 1; do
  echo "== $sc L2_PERSIST=$v"
  MRT_L2_VERBOSE=1 MRT_L2_PERSIST=$v timeout 900 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*//'
done; done; done 2>&1 | tee gpurun_out/l2_persist_ac.log
