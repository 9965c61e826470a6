#!/bin/bash
# round-1 GPU pass AV: vote tuning of the production trace kernel re-swept on the final kernel and tree
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32"; do
for lm in 6 8 10 12; do for tr in 4 5 6 7 8; do for rf in 24; do
  echo "== $sc LEAFMIN=$lm TRIRATIO=$tr REFILL=$rf"
  MRT_TRACE_LEAFMIN=$lm MRT_TRACE_TRIRATIO=$tr MRT_TRACE_REFILL=$rf MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done; done
for rf in 20 22 26 28; do
  echo "== $sc LEAFMIN=8 TRIRATIO=6 REFILL=$rf"
  MRT_TRACE_REFILL=$rf MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done
done 2>&1 | tee gpurun_out/vote_av.log
