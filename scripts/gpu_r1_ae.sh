#!/bin/bash
# round-1 GPU pass AE: trace kernel with the next node chosen ahead (MRT_TRACE_AHEAD: 1 form only, 2 L1 prefetch, 3 L2 prefetch)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x -k "ahead" 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/pytest_gpu_ae.log
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 0 1 2 3; do
  echo "== $sc AHEAD=$v"
  MRT_TRACE_AHEAD=$v MRT_WAVE_LOG=/tmp/w.csv timeout 900 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*//'
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done 2>&1 | tee gpurun_out/ahead_ae.log
