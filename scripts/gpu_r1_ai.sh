#!/bin/bash
# round-1 GPU pass AI: triangle-queue trace kernel, second version (ballot prefix push, cheaper retire, wait trigger)
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
MRT_TRACE_TQ=32 timeout 60 $R --scene cornell --size 256x256 --frames 4 --seed 1 --pipeline wavefront 2>&1 | cut -c1-200 || { echo "SMALL RUN FAILED/HUNG"; exit 1; }
( timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x -k "tq" 2>&1 | tail -15 ) 2>&1 | tee gpurun_out/pytest_gpu_ai.log
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32"; do
for v in "0 8" "16 8" "24 4" "24 8" "24 33" "32 8" "32 33"; do
  set -- $v
  echo "== $sc TQ=$1 WAIT=$2"
  MRT_TRACE_COUNT=1 MRT_TRACE_TQ=$1 MRT_TRACE_TQWAIT=$2 timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*"node_tests"/"node_tests"/' | cut -c1-300
  MRT_TRACE_TQ=$1 MRT_TRACE_TQWAIT=$2 MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/tq_ai.log
