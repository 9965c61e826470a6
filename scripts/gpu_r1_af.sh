#!/bin/bash
# round-1 GPU pass AF: trace kernel with the cold per-ray state parked in shared memory (MRT_TRACE_PARK=1) at 8 / 9 / 10 blocks per SM
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x -k "park" 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/pytest_gpu_af.log
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in "0 8" "1 8" "1 9" "1 10"; do
  set -- $v
  echo "== $sc PARK=$1 MINBLOCKS=$2"
  MRT_TRACE_PARK=$1 MRT_TRACE_MINBLOCKS=$2 MRT_WAVE_LOG=/tmp/w.csv timeout 900 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*//'
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done 2>&1 | tee gpurun_out/park_af.log
