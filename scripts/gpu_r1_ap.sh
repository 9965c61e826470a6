#!/bin/bash
# round-1 GPU pass AP: SAH bins per axis of the binary builder (MRT_BVH_BINS), tests per ray and trace time
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 16 8 32 64; do
  echo "== $sc BINS=$v"
  MRT_BVH_BINS=$v MRT_TRACE_COUNT=1 timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*"bvh_build_ms"/"bvh_build_ms"/' | cut -c1-400
  MRT_BVH_BINS=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/bins_ap.log
