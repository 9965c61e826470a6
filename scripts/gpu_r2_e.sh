#!/bin/bash
# round-2 GPU pass E: re-sweep of the trace kernel's tuning on the loop-free leaf expansion (experiments build): blocks per SM,
# vote thresholds, refill threshold, slot groups; plus the launch list of the bench command under the host-loop schedule and
# the new schedule / closed-form GPU tests.
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render_exp
run() { echo "== $*"; env "$@" timeout 300 $R $S --pipeline wavefront 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-14s mrays %8.1f device_ms %8.2f trace_ms %8.2f'%(d['scene'],d['mrays_per_s'],d['device_ms'],d['profile_trace_ms']))"; }
{
for S in "--scene spheres-1m --size 1920x1080 --seed 3 --frames 64" "--scene many-lights --size 1920x1080 --seed 5 --frames 64"; do
  run MRT_WAVE_LOG=/dev/null
  run MRT_WAVE_LOG=/dev/null
  for mb in 9 10; do run MRT_WAVE_LOG=/dev/null MRT_TRACE_MINBLOCKS=$mb; done
  for lm in 8 12 14; do run MRT_WAVE_LOG=/dev/null MRT_TRACE_LEAFMIN=$lm; done
  for tr in 5 6 8 9; do run MRT_WAVE_LOG=/dev/null MRT_TRACE_TRIRATIO=$tr; done
  for rf in 20 22 26 28; do run MRT_WAVE_LOG=/dev/null MRT_TRACE_REFILL=$rf; done
  run MRT_WF_GRAPH=0
  run MRT_WF_GRAPH=0 MRT_WF_SPLIT=2
  run MRT_WF_GRAPH=1
done
} 2>&1 | tee gpurun_out/e_sweep.log
( timeout 900 python -m pytest tests/test_branch_closed_forms.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/e_pytest.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ) | tee gpurun_out/e_smoke.log
MRT_WF_GRAPH=0 timeout 600 python bench.py --steps 1 --warmup 3 --frames 32 --no-cpu-baseline --no-parity > gpurun_out/e_plain_bench.log 2>&1 && \
MRT_WF_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 800 -c 400 --csv --log-file gpurun_out/e_launches_bench.csv \
    python bench.py --steps 1 --warmup 3 --frames 32 --no-cpu-baseline --no-parity > gpurun_out/e_ncu_bench.log 2>&1
wc -l gpurun_out/e_launches_bench.csv
