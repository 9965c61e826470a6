#!/bin/bash
# round-1 GPU pass A: parity tests, occupancy variants, bench, ncu launch list + full capture of the fused kernel
# (each ncu run follows the same plain command exiting 0)
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv
nproc
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu_a.log
for v in 2 3 4; do MRT_FUSED_MINBLOCKS=$v timeout 300 $R --scene spheres-1m --size 1920x1080 --frames 16 --seed 3; done 2>&1 | tee gpurun_out/variants_a.log
timeout 300 $R --scene spheres-1m --size 1920x1080 --frames 4 --seed 3 --pipeline staged 2>&1 | tee -a gpurun_out/variants_a.log
timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 2>&1 | tee -a gpurun_out/variants_a.log
timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 2>&1 | tee -a gpurun_out/variants_a.log
timeout 600 python bench.py --steps 2 --warmup 3 --frames 32 > gpurun_out/bench_r1_a.json 2> gpurun_out/bench_r1_a.err; tail -c 3000 gpurun_out/bench_r1_a.json; tail -5 gpurun_out/bench_r1_a.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_a_ref.json 2> gpurun_out/bench_r1_a_ref.err; tail -c 1500 gpurun_out/bench_r1_a_ref.json
# launch list of the reference-shaped sequence (1 frame = 34 dispatches) and of the fused kernel
timeout 300 $R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/plain_staged.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_staged.csv \
    $R --scene spheres-1m --size 1920x1080 --frames 1 --seed 3 --pipeline staged > gpurun_out/ncu_staged.log 2>&1
timeout 300 $R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 > gpurun_out/plain_fused.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -c 1 -o gpurun_out/prof_fused_a \
    $R --scene spheres-1m --size 1920x1080 --frames 2 --seed 3 > gpurun_out/ncu_fused.log 2>&1
ls -la gpurun_out
