#!/bin/bash
# A/B of several builds of the library on one box: metal-renderer_b200/lib_ab/<name>.so (built locally, not tracked), given
# as arguments; prints device time and summed trace-kernel time per build and scene, REPS repetitions (default 2)
cd metal-renderer_b200
cp lib/libmrt_b200.so /tmp/keep.so
for rep in $(seq 1 ${REPS:-2}); do
for v in "$@"; do
  cp lib_ab/$v.so lib/libmrt_b200.so
  for sc in "spheres-1m --size 1920x1080 --frames 32" "cornell-water --size 1920x1080 --frames 32" "many-lights --size 1920x1080 --frames 32" "spheres-10m --size 3840x2160 --frames 8"; do
  MRT_WAVE_LOG=/tmp/w.csv ./bin/mrt_render --scene $sc --seed 3 --pipeline wavefront > /tmp/o.json
  python - "$v" "$sc" <<'PY'
import json, sys, csv
d = json.load(open('/tmp/o.json'))
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print(sys.argv[1], sys.argv[2].split()[0], "device_ms %.2f trace_ms %.2f rest_ms %.2f mrays %.1f" % (d['device_ms'], t, d['device_ms'] - t, d['mrays_per_s']))
PY
  done
done
done
cp /tmp/keep.so lib/libmrt_b200.so
