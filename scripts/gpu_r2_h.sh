#!/bin/bash
# round-2 GPU pass H: the record at HEAD — full GPU test-suite, smoke, the bench line as the driver runs it, the reference arm,
# configs 4 and 5 at their full samples per pixel, and the ncu summary of a STEADY-STATE wave of the 10 M-triangle scene.
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -5 ) | tee gpurun_out/h_pytest.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ) | tee gpurun_out/h_smoke.log
R=./metal-renderer_b200/bin/mrt_render
wl=spheres-10m-4k; args="--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"; skip=14
MRT_WF_GRAPH=0 MRT_WAVE_LOG=gpurun_out/h_waves_$wl.csv timeout 600 $R $args --pipeline wavefront > gpurun_out/h_plain_$wl.log 2>&1
rays=$(awk -F, -v w=$skip 'NR>1 && $1==w {print $2+$3}' gpurun_out/h_waves_$wl.csv)
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s $skip -c 1 -f -o gpurun_out/h_trace_$wl $R $args --pipeline wavefront > gpurun_out/h_ncu_trace_$wl.log 2>&1
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s $((skip+1)) -c 1 -f -o gpurun_out/h_shade_$wl $R $args --pipeline wavefront > gpurun_out/h_ncu_shade_$wl.log 2>&1
python scripts/ncu_collect.py gpurun_out/h_trace_$wl.ncu-rep trace $wl ${rays:-0} "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_trace -s $skip -c 1" | sed -n 4,12p
python scripts/ncu_collect.py gpurun_out/h_shade_$wl.ncu-rep shade $wl 0 "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_shade -s $((skip+1)) -c 1" | sed -n 5,6p
cp profiles/ncu_*_$wl.txt gpurun_out/
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; tail -c 1200 gpurun_out/h_bench.json; tail -3 gpurun_out/h_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/h_bench_ref.json 2> gpurun_out/h_bench_ref.err; tail -c 500 gpurun_out/h_bench_ref.json
timeout 900 python bench.py --workload many-lights-1080p --steps 2 --no-cpu-baseline > gpurun_out/h_bench_cfg5_512spp.json 2>> gpurun_out/h_bench.err; tail -c 500 gpurun_out/h_bench_cfg5_512spp.json
timeout 1200 python bench.py --workload spheres-10m-4k --steps 2 --no-cpu-baseline > gpurun_out/h_bench_cfg4_256spp.json 2>> gpurun_out/h_bench.err; tail -c 500 gpurun_out/h_bench_cfg4_256spp.json
