#!/bin/bash
# round-1 GPU pass AR: ground truth for the tree selection — fixed SAH bin counts on the default workload and many-lights
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for sc in "spheres-1m --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32"; do
for v in 6 10 12 14 16 18 20 24 28 40 48 56 64; do
  echo "== $sc BINS=$v"
  MRT_BVH_BINS=$v MRT_TRACE_COUNT=1 timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/.*"node_tests"/"node_tests"/' | cut -c1-100
  MRT_BVH_BINS=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/bins_ar.log
