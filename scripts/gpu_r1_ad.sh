#!/bin/bash
# round-1 GPU pass AD: GPU tests of the committed build + sensitivity of the trace kernel to resident warps
# (MRT_TRACE_MINBLOCKS = blocks of 128 threads per SM; 8 is the default)
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/pytest_gpu_ad.log
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for mb in 5 6 8; do
  echo "== minblocks=$mb"
  MRT_TRACE_MINBLOCKS=$mb MRT_WAVE_LOG=/tmp/w.csv timeout 900 $R --scene spheres-1m --size 1920x1080 --frames 32 --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*//'
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/minblocks_ad.log
