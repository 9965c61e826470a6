#!/bin/bash
# round-1 GPU pass AW: vote tuning candidates on all four scenes, two repetitions
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for rep in 1 2; do
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
for v in "8 6" "10 7" "10 8" "12 8"; do
  set -- $v
  echo "== $sc LEAFMIN=$1 TRIRATIO=$2"
  MRT_TRACE_LEAFMIN=$1 MRT_TRACE_TRIRATIO=$2 MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done 2>&1 | tee gpurun_out/vote_aw.log
