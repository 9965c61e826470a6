#!/bin/bash
# round-2 GPU pass A: new parity tests (BASELINE sizes vs oracle, sharded step on one GPU), the WHILE-graph schedule and the
# ray-ordering experiment (MRT_WF_SORT) A/B, ncu baselines of trace + shade on the headline workload.
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader | tee gpurun_out/a_gpu.txt
nproc | tee -a gpurun_out/a_gpu.txt
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_fuzz.py -x -q 2>&1 | tail -15 ) | tee gpurun_out/a_pytest1.log
R=./metal-renderer_b200/bin/mrt_render_exp
S1="--scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront"
run() { echo "== $*"; env "$@" timeout 300 $R $S 2>&1 | tail -2; }
{
S="$S1"
run MRT_WF_GRAPH=1
run MRT_WF_GRAPH=0
run MRT_WF_GRAPH=1
for b in 1 2 3; do for m in 0 1; do run MRT_WF_SORT=$b MRT_WF_SORT_MODE=$m; done; done
run MRT_WAVE_LOG=gpurun_out/a_waves_base.csv
run MRT_WAVE_LOG=gpurun_out/a_waves_sort2.csv MRT_WF_SORT=2
run MRT_WAVE_LOG=gpurun_out/a_waves_sort3.csv MRT_WF_SORT=3
run MRT_WF_SORT=3 MRT_TRACE_TQ=32
run MRT_WF_SORT=3 MRT_TRACE_LEAFMIN=12 MRT_TRACE_TRIRATIO=8
run MRT_WF_SORT=3 MRT_TRACE_REFILL=28
run MRT_WF_SORT=3 MRT_TRACE_REFILL=20
S="--scene cornell-water --size 1920x1080 --seed 2 --frames 64 --pipeline wavefront"
run MRT_WF_SORT=0; run MRT_WF_SORT=2; run MRT_WF_SORT=3
S="--scene many-lights --size 1920x1080 --seed 5 --frames 64 --pipeline wavefront"
run MRT_WF_SORT=0; run MRT_WF_SORT=2; run MRT_WF_SORT=3
S="--scene spheres-10m --size 3840x2160 --seed 4 --frames 8 --pipeline wavefront"
run MRT_WF_SORT=0; run MRT_WF_SORT=3
} 2>&1 | tee gpurun_out/a_sweep.log
( timeout 1500 python -m pytest tests/test_gpu_baseline_sizes.py -x -q 2>&1 | tail -15 ) | tee gpurun_out/a_pytest2.log
# ncu: steady-state trace and shade launches, baseline and sorted
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
for v in 0 3; do
  MRT_WF_GRAPH=0 MRT_WF_SORT=$v timeout 300 $R $S > gpurun_out/a_plain_$v.log 2>&1 && \
  MRT_WF_GRAPH=0 MRT_WF_SORT=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -f -o gpurun_out/a_trace_sort$v $R $S > gpurun_out/a_ncu_trace_$v.log 2>&1
  MRT_WF_GRAPH=0 MRT_WF_SORT=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 30 -c 1 -f -o gpurun_out/a_shade_sort$v $R $S > gpurun_out/a_ncu_shade_$v.log 2>&1
done
ls -la gpurun_out | tail -20
( timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_display.py tests/test_mitsuba_sanity.py -x -q -m gpu 2>&1 | tail -15 ) | tee gpurun_out/a_pytest3.log
( timeout 600 python bench.py --steps 1 --warmup 3 --frames 16 --no-cpu-baseline 2> gpurun_out/a_bench_smoke.err | tail -c 6000 ) | tee gpurun_out/a_bench_smoke.json
tail -5 gpurun_out/a_bench_smoke.err
