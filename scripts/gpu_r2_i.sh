#!/bin/bash
# round-2 GPU pass I: the no-alloc entry point's test ALONE in a fresh process, then the whole GPU suite (the same test after
# all the others), smoke, and the record at HEAD that pass H took (lost with the machine): bench as the driver runs it, the
# reference arm, configs 4 and 5 at their full samples per pixel, ncu of a steady-state wave of the 10 M-triangle scene.
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4 ) | tee gpurun_out/i_pytest_noalloc_alone.log
( timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) | tee gpurun_out/i_pytest.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ) | tee gpurun_out/i_smoke.log
R=./metal-renderer_b200/bin/mrt_render
wl=spheres-10m-4k; args="--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"; skip=14
MRT_WF_GRAPH=0 MRT_WAVE_LOG=gpurun_out/i_waves_$wl.csv timeout 600 $R $args --pipeline wavefront > gpurun_out/i_plain_$wl.log 2>&1
rays=$(awk -F, -v w=$skip 'NR>1 && $1==w {print $2+$3}' gpurun_out/i_waves_$wl.csv)
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s $skip -c 1 -f -o gpurun_out/i_trace_$wl $R $args --pipeline wavefront > gpurun_out/i_ncu_trace_$wl.log 2>&1
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s $((skip+1)) -c 1 -f -o gpurun_out/i_shade_$wl $R $args --pipeline wavefront > gpurun_out/i_ncu_shade_$wl.log 2>&1
python scripts/ncu_collect.py gpurun_out/i_trace_$wl.ncu-rep trace $wl ${rays:-0} "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_trace -s $skip -c 1" | sed -n 4,12p
python scripts/ncu_collect.py gpurun_out/i_shade_$wl.ncu-rep shade $wl 0 "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_shade -s $((skip+1)) -c 1" | sed -n 5,6p
cp profiles/ncu_*_$wl.txt gpurun_out/
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; tail -c 1200 gpurun_out/i_bench.json; tail -3 gpurun_out/i_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/i_bench_ref.json 2> gpurun_out/i_bench_ref.err; tail -c 500 gpurun_out/i_bench_ref.json
timeout 900 python bench.py --workload many-lights-1080p --steps 2 --no-cpu-baseline > gpurun_out/i_bench_cfg5_512spp.json 2>> gpurun_out/i_bench.err; tail -c 500 gpurun_out/i_bench_cfg5_512spp.json
timeout 1200 python bench.py --workload spheres-10m-4k --steps 2 --no-cpu-baseline > gpurun_out/i_bench_cfg4_256spp.json 2>> gpurun_out/i_bench.err; tail -c 500 gpurun_out/i_bench_cfg4_256spp.json
