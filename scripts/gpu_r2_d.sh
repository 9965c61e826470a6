#!/bin/bash
# round-2 GPU pass D: ncu --set full of one steady-state trace launch and one shade launch per workload (summaries written to
# profiles/ncu_{trace,shade}_<workload>.txt on the box, so that the bench lines below parse them), bench lines of every config,
# the reference arm, and the ncu launch list of the bench command.
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
cap() {   # workload scene-args skip
  local wl=$1 args=$2 skip=$3
  MRT_WF_GRAPH=0 MRT_WAVE_LOG=gpurun_out/d_waves_$wl.csv timeout 600 $R $args --pipeline wavefront > gpurun_out/d_plain_$wl.log 2>&1 || { echo "plain run failed for $wl"; return; }
  local rays=$(awk -F, -v w=$skip 'NR>1 && $1==w {print $2+$3}' gpurun_out/d_waves_$wl.csv)
  MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s $skip -c 1 -f -o gpurun_out/d_trace_$wl $R $args --pipeline wavefront > gpurun_out/d_ncu_trace_$wl.log 2>&1
  MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s $((skip+1)) -c 1 -f -o gpurun_out/d_shade_$wl $R $args --pipeline wavefront > gpurun_out/d_ncu_shade_$wl.log 2>&1
  python scripts/ncu_collect.py gpurun_out/d_trace_$wl.ncu-rep trace $wl ${rays:-0} "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_trace -s $skip -c 1" | sed -n 5,12p
  python scripts/ncu_collect.py gpurun_out/d_shade_$wl.ncu-rep shade $wl 0 "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_shade -s $((skip+1)) -c 1" | sed -n 5,6p
}
cap spheres-1m-1080p "--scene spheres-1m --size 1920x1080 --seed 3 --frames 8" 30
cap cornell-water-1080p "--scene cornell-water --size 1920x1080 --seed 2 --frames 8" 30
cap many-lights-1080p "--scene many-lights --size 1920x1080 --seed 5 --frames 8" 30
cap spheres-10m-4k "--scene spheres-10m --size 3840x2160 --seed 4 --frames 4" 20
cp profiles/ncu_*.txt gpurun_out/ 2>/dev/null
timeout 900 python bench.py > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; tail -c 1500 gpurun_out/d_bench.json; tail -3 gpurun_out/d_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/d_bench_ref.json 2> gpurun_out/d_bench_ref.err; tail -c 800 gpurun_out/d_bench_ref.json
timeout 900 python bench.py --workload cornell-water-1080p --steps 2 --no-cpu-baseline > gpurun_out/d_bench_cfg2.json 2>> gpurun_out/d_bench.err; tail -c 600 gpurun_out/d_bench_cfg2.json
timeout 900 python bench.py --workload media-water-plastic-1080p --steps 2 --no-cpu-baseline > gpurun_out/d_bench_cfg2_media.json 2>> gpurun_out/d_bench.err; tail -c 600 gpurun_out/d_bench_cfg2_media.json
timeout 900 python bench.py --workload many-lights-1080p --frames 64 --steps 2 --no-cpu-baseline > gpurun_out/d_bench_cfg5.json 2>> gpurun_out/d_bench.err; tail -c 600 gpurun_out/d_bench_cfg5.json
timeout 900 python bench.py --workload spheres-10m-4k --frames 16 --steps 2 --no-cpu-baseline > gpurun_out/d_bench_cfg4.json 2>> gpurun_out/d_bench.err; tail -c 600 gpurun_out/d_bench_cfg4.json
timeout 600 python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline --no-parity > gpurun_out/d_plain_bench.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/d_launches_bench.csv \
    python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline --no-parity > gpurun_out/d_ncu_bench.log 2>&1
ls -la gpurun_out | tail -30
