#!/bin/bash
# round-2 GPU pass U: the record at HEAD after the optics extension / peer reduction went in (default kernels' SASS unchanged):
# the no-alloc test alone, the whole GPU suite, smoke, ncu of a steady-state trace and shade launch of the headline workload,
# bench as the driver runs it, the reference arm, the launch list of the bench command.
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3 ) | tee gpurun_out/u_pytest_noalloc_alone.log
( timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) | tee gpurun_out/u_pytest.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ) | tee gpurun_out/u_smoke.log
R=./metal-renderer_b200/bin/mrt_render
wl=spheres-1m-1080p; args="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8"; skip=30
MRT_WF_GRAPH=0 MRT_WAVE_LOG=gpurun_out/u_waves_$wl.csv timeout 600 $R $args --pipeline wavefront > gpurun_out/u_plain_$wl.log 2>&1
rays=$(awk -F, -v w=$skip 'NR>1 && $1==w {print $2+$3}' gpurun_out/u_waves_$wl.csv)
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s $skip -c 1 -f -o gpurun_out/u_trace_$wl $R $args --pipeline wavefront > gpurun_out/u_ncu_trace_$wl.log 2>&1
MRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s $((skip+1)) -c 1 -f -o gpurun_out/u_shade_$wl $R $args --pipeline wavefront > gpurun_out/u_ncu_shade_$wl.log 2>&1
python scripts/ncu_collect.py gpurun_out/u_trace_$wl.ncu-rep trace $wl ${rays:-0} "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_trace -s $skip -c 1" | sed -n 4,12p
python scripts/ncu_collect.py gpurun_out/u_shade_$wl.ncu-rep shade $wl 0 "MRT_WF_GRAPH=0 mrt_render $args --pipeline wavefront; ncu -k regex:k_wf_shade -s $((skip+1)) -c 1" | sed -n 5,6p
cp profiles/ncu_*_$wl.txt gpurun_out/
rm -f gpurun_out/u_*.ncu-rep
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/u_bench.json 2> gpurun_out/u_bench.err; tail -c 1200 gpurun_out/u_bench.json; tail -3 gpurun_out/u_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/u_bench_ref.json 2> gpurun_out/u_bench_ref.err; tail -c 500 gpurun_out/u_bench_ref.json
timeout 600 python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline --no-parity > gpurun_out/u_plain_bench.log 2>&1 && \
MRT_WF_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/u_launches_bench.csv \
    python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline --no-parity > gpurun_out/u_ncu_bench.log 2>&1
ls -la gpurun_out | tail -20
