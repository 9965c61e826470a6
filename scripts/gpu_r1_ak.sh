#!/bin/bash
# round-1 GPU pass AK/AL: triangle-queue trace kernel, later versions
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
MRT_TRACE_TQ=32 timeout 60 $R --scene cornell --size 256x256 --frames 4 --seed 1 --pipeline wavefront 2>&1 | cut -c1-200 || { echo "SMALL RUN FAILED/HUNG"; exit 1; }
( timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x -k "tq" 2>&1 | tail -15 ) 2>&1 | tee gpurun_out/pytest_gpu_ak.log
for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32"; do
for v in "0 24" "12 24" "16 24" "24 24" "32 24"; do
  set -- $v
  echo "== $sc TQ=$1 REFILL=$2"
  MRT_TRACE_TQ=$1 MRT_TRACE_REFILL=$2 MRT_WAVE_LOG=/tmp/w.csv timeout 300 $R --scene $sc --seed 3 --pipeline wavefront > /dev/null 2>&1
  python - <<'PY'
import csv
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print("trace_ms %.2f" % t)
PY
done; done 2>&1 | tee gpurun_out/tq_ak.log
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
MRT_TRACE_TQ=24 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -f -o gpurun_out/prof_al_tq24 $R $S > gpurun_out/ncu_al_tq24.log 2>&1
