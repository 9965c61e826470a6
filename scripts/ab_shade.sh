#!/bin/bash
# A/B of two builds of the library on one box: device time of a 64-frame render and the trace kernel's share of it;
# variants = metal-renderer_b200/lib_ab/*.so (built locally, not tracked)
cd metal-renderer_b200
for rep in 1 2 3; do
for v in old new; do
  cp lib_ab/$v.so lib/libmrt_b200.so
  for sc in spheres-1m cornell-water; do
  MRT_WAVE_LOG=/tmp/w.csv ./bin/mrt_render --scene $sc --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront > /tmp/o.json
  python - "$v" "$sc" <<'PY'
import json, sys, csv
d = json.load(open('/tmp/o.json'))
t = sum(float(r['trace_ms']) for r in csv.DictReader(open('/tmp/w.csv')))
print(sys.argv[1], sys.argv[2], "device_ms %.2f trace_ms %.2f rest_ms %.2f mrays %.1f" % (d['device_ms'], t, d['device_ms'] - t, d['mrays_per_s']))
PY
  done
done
done
cp lib_ab/new.so lib/libmrt_b200.so
