#!/bin/bash
# round-1 GPU pass AY: triangle cost of the wide-tree collapse (MRT_BVH8_CPRIM) re-swept on the final kernel
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 0.6 0.8 1.0 1.25 1.5 2.0 3.0; do
  echo "== $sc CPRIM=$v"
  MRT_BVH8_CPRIM=$v MRT_TRACE_COUNT=1 timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/.*"node_tests"/"node_tests"/' | cut -c1-100
  MRT_BVH8_CPRIM=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/cprim_ay.log
