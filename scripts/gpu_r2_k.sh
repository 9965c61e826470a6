#!/bin/bash
# round-2 GPU pass K: host BVH build made parallel where it was serial (level-parallel emit, collapse over subtrees, tables
# not zero-filled, sampling beside the next candidate's build, binary tree beside the wide one) — same trees bit for bit.
# Confirms parity on the GPU and measures set-up time on the box's host cores.
mkdir -p gpurun_out
nproc | tee gpurun_out/k_cores.txt
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/k_pytest.log
R=./metal-renderer_b200/bin/mrt_render
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 16" "--scene spheres-10m --size 3840x2160 --seed 4 --frames 4" "--scene cornell-water --size 1920x1080 --seed 2 --frames 16"; do
  for rep in 1 2; do echo "== $S"; MRT_BVH_VERBOSE=1 timeout 600 $R $S --pipeline wavefront 2>&1 | grep -E "records|mrays" | sed -E 's/.*"mrays_per_s": ([0-9.]+).*"bvh_build_ms": ([0-9.]+).*"bvh8_build_ms": ([0-9.]+).*/mrays \1 bvh_build_ms \2 bvh8_build_ms \3/'; done
done
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128 --depth 8"
timeout 300 $R $S --devices 0 2>/dev/null | head -1 | cut -c1-200
} 2>&1 | tee gpurun_out/k_build_times.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/k_bench.json').read().strip().splitlines()[-1]); print('value',d['value'],'e2e',d['e2e']['value'],'setup',d['setup_ms'],'bvh',d['bvh'],'parity',d['parity']['pass'])" | tee gpurun_out/k_bench_summary.txt
