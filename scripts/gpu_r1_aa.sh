#!/bin/bash
# round-1 GPU pass AA / AZ: re-validation of HEAD on a fresh box (tests, smoke, default bench)
set -x
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 ) 2>&1 | tee gpurun_out/pytest_gpu_az.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke_az.log
timeout 900 python bench.py > gpurun_out/bench_r1_az.json 2> gpurun_out/bench_r1_az.err; tail -c 1500 gpurun_out/bench_r1_az.json; tail -5 gpurun_out/bench_r1_az.err
