#!/bin/bash
# round-1 GPU pass G: full tests + bench (ours, reference arm) + ncu launch list of the bench command
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_g.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke_g.log
timeout 900 python bench.py > gpurun_out/bench_r1_g.json 2> gpurun_out/bench_r1_g.err; tail -c 4000 gpurun_out/bench_r1_g.json; tail -5 gpurun_out/bench_r1_g.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r1_g_ref.json 2> gpurun_out/bench_r1_g_ref.err; tail -c 1500 gpurun_out/bench_r1_g_ref.json
timeout 900 python bench.py --workload cornell-water-1080p --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_g_cfg2.json 2>> gpurun_out/bench_r1_g.err; tail -c 2500 gpurun_out/bench_r1_g_cfg2.json
timeout 600 python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/launches_bench_g.csv \
    python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
ls -la gpurun_out
