#!/bin/bash
# round-1 GPU pass AT: bench.py after the stdout change — N=1 (ours, reference arm) as the driver runs it
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_r1_at.json 2> gpurun_out/bench_r1_at.err; echo "rc=$? lines=$(wc -l < gpurun_out/bench_r1_at.json)"; head -c 300 gpurun_out/bench_r1_at.json; echo; tail -3 gpurun_out/bench_r1_at.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r1_at_ref.json 2> gpurun_out/bench_r1_at_ref.err; echo "rc=$? lines=$(wc -l < gpurun_out/bench_r1_at_ref.json)"; head -c 300 gpurun_out/bench_r1_at_ref.json; echo
timeout 900 python bench.py --workload cornell-water-1080p --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_at_cfg2.json 2>> gpurun_out/bench_r1_at.err; head -c 200 gpurun_out/bench_r1_at_cfg2.json; echo
timeout 900 python bench.py --workload spheres-10m-4k --frames 16 --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_at_cfg4.json 2>> gpurun_out/bench_r1_at.err; head -c 200 gpurun_out/bench_r1_at_cfg4.json; echo
timeout 900 python bench.py --workload many-lights-1080p --frames 64 --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_at_cfg5.json 2>> gpurun_out/bench_r1_at.err; head -c 200 gpurun_out/bench_r1_at_cfg5.json; echo
