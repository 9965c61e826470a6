#!/bin/bash
# round-2 GPU pass L: A/B of the load order in intersection_handler (noise sample and light selection requested before the
# shading record is waited for) against the previous commit's build: shade time = render time - event-timed trace time.
mkdir -p gpurun_out
NEW=./metal-renderer_b200/bin/mrt_render; OLD=./metal-renderer_b200/prev/bin/mrt_render
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 64" "--scene many-lights --size 1920x1080 --seed 5 --frames 64" \
         "--scene cornell-water --size 1920x1080 --seed 2 --frames 64" "--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"; do
  for rep in 1 2 3; do for B in $OLD $NEW; do echo "== $B $S"; MRT_WAVE_LOG=/dev/null timeout 600 $B $S --pipeline wavefront | tail -1; done; done
  for B in $OLD $NEW; do echo "== graph $B $S"; timeout 600 $B $S --pipeline wavefront | tail -1; done
done
for P in fused staged; do for B in $OLD $NEW; do echo "== $P $B"; timeout 600 $B --scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline $P | tail -1; done; done
} 2>&1 | tee gpurun_out/l_ab.log
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_branch_closed_forms.py tests/test_gpu_fuzz.py tests/test_gpu_baseline_sizes.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/l_pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/l_bench.json').read().strip().splitlines()[-1]); print('value',d['value'],'e2e',d['e2e']['value'],'setup',d['setup_ms'],'bvh',d['bvh']['build_ms'],d['bvh']['wide_build_ms'],'parity',d['parity']['pass'])" | tee gpurun_out/l_bench_summary.txt
