#!/bin/bash
# round-1 GPU pass AQ: wide tree chosen among SAH-bin candidates by sampled traversal cost — full tests + all workloads
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/pytest_gpu_aq.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke_aq.log
R=./metal-renderer_b200/bin/mrt_render
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in 16 16,64; do
  echo "== $sc BINS=$v"
  MRT_BVH_VERBOSE=1 MRT_BVH_BINS=$v MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront 2>&1 | sed -E 's/"triangles.*"device_ms"/"device_ms"/; s/"primary_rays.*//' | cut -c1-200
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/bins_aq.log
timeout 900 python bench.py > gpurun_out/bench_r1_aq.json 2> gpurun_out/bench_r1_aq.err; tail -c 600 gpurun_out/bench_r1_aq.json
