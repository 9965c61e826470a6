#!/bin/bash
# round-1 GPU pass AN: shared-memory carve-out hint of the production trace kernel (more L1 for the tree)
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in -1 0 10 25 50; do
  echo "== $sc CARVEOUT=$v"
  MRT_TRACE_CARVEOUT=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done 2>&1 | tee gpurun_out/carveout_an.log
