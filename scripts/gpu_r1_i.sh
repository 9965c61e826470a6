#!/bin/bash
# round-1 GPU pass I (per-region work counters + stealing): parity, throughput, bench
set -x
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu_i.log
S="--scene spheres-1m --size 1920x1080 --seed 3"
{ timeout 300 $R $S --frames 16 --pipeline wavefront
  timeout 300 $R $S --frames 128 --pipeline wavefront
  timeout 300 $R --scene cornell-water --size 1920x1080 --frames 64 --seed 2 --pipeline wavefront
  timeout 300 $R --scene many-lights --size 1920x1080 --frames 16 --seed 5 --pipeline wavefront
  timeout 600 $R --scene spheres-10m --size 3840x2160 --frames 4 --seed 4 --pipeline wavefront
} 2>&1 | tee gpurun_out/sweep_i.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_r1_i.json 2> gpurun_out/bench_r1_i.err; tail -c 2500 gpurun_out/bench_r1_i.json; tail -3 gpurun_out/bench_r1_i.err
