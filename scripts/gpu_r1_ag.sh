#!/bin/bash
# round-1 GPU pass AG: trace kernel with a warp-wide triangle queue (MRT_TRACE_TQ = pairs that trigger a triangle step)
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
MRT_TRACE_TQ=32 timeout 60 $R --scene cornell --size 256x256 --frames 4 --seed 1 --pipeline wavefront 2>&1 | cut -c1-300 || { echo "SMALL RUN FAILED/HUNG"; exit 1; }
( timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x -k "tq" 2>&1 | tail -15 ) 2>&1 | tee gpurun_out/pytest_gpu_ag.log
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 0 16 24 32 48; do
  echo "== $sc TQ=$v"
  MRT_TRACE_COUNT=1 MRT_TRACE_TQ=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*"node_tests"/"node_tests"/' | cut -c1-300
  MRT_TRACE_TQ=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done 2>&1 | tee gpurun_out/tq_ag.log
