#!/bin/bash
# tests, variants, bench, ncu launch list + full capture (each ncu run follows the same plain command exiting 0)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for v in 2 3 4; do MRT_FUSED_MINBLOCKS=$v $R --scene spheres-1m --size 1920x1080 --frames 16 --seed 3; done
$R --scene spheres-1m --size 1920x1080 --frames 16 --seed 3 --pipeline staged
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r1_a.json 2> gpurun_out/bench_r1_a.err; tail -c 3000 gpurun_out/bench_r1_a.json; tail -5 gpurun_out/bench_r1_a.err
# launch list of the reference-shaped sequence (1 frame = 34 dispatches) and of the fused kernel
$R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/plain_staged.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_staged.csv \
    $R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/ncu_staged.log 2>&1
$R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 > gpurun_out/plain_fused.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_fused -c 1 -o gpurun_out/prof_fused \
    $R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 > gpurun_out/ncu_fused.log 2>&1
$R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/plain_staged2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_intersect -s 4 -c 2 -o gpurun_out/prof_intersect \
    $R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/ncu_intersect.log 2>&1
ls -la gpurun_out
