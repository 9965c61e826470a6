#!/bin/bash
# round-1 GPU pass AA: re-validation of HEAD on a fresh box (tests, smoke, default bench)
set -x
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 ) 2>&1 | tee gpurun_out/pytest_gpu_aa.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke_aa.log
timeout 900 python bench.py > gpurun_out/bench_r1_aa.json 2> gpurun_out/bench_r1_aa.err; tail -c 1500 gpurun_out/bench_r1_aa.json; tail -5 gpurun_out/bench_r1_aa.err
