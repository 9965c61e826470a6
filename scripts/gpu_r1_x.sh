#!/bin/bash
# round-1 GPU pass X: full tests + smoke + bench (ours, reference arm, config 2) + ncu launch list of the bench command
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_x.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke_x.log
timeout 900 python bench.py > gpurun_out/bench_r1_x.json 2> gpurun_out/bench_r1_x.err; tail -c 4000 gpurun_out/bench_r1_x.json; tail -5 gpurun_out/bench_r1_x.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r1_x_ref.json 2> gpurun_out/bench_r1_x_ref.err; tail -c 1500 gpurun_out/bench_r1_x_ref.json
timeout 900 python bench.py --workload cornell-water-1080p --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_x_cfg2.json 2>> gpurun_out/bench_r1_x.err; tail -c 2500 gpurun_out/bench_r1_x_cfg2.json
timeout 900 python bench.py --workload many-lights-1080p --frames 64 --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_x_cfg5.json 2>> gpurun_out/bench_r1_x.err; tail -c 2500 gpurun_out/bench_r1_x_cfg5.json
timeout 900 python bench.py --workload spheres-10m-4k --frames 16 --steps 2 --no-cpu-baseline > gpurun_out/bench_r1_x_cfg4.json 2>> gpurun_out/bench_r1_x.err; tail -c 2500 gpurun_out/bench_r1_x_cfg4.json
timeout 600 python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/plain_bench_x.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/launches_bench_x.csv \
    python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/ncu_bench_x.log 2>&1
R=./metal-renderer_b200/bin/mrt_render
for rf in 20 24 28; do echo "refill=$rf"; MRT_TRACE_REFILL=$rf timeout 300 $R --scene spheres-1m --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront; done 2>&1 | grep -v "^+" | tee gpurun_out/sweep_x.log
ls -la gpurun_out | tail -12
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront"
timeout 300 $R $S > gpurun_out/plain_x.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -o gpurun_out/prof_wf_x_trace $R $S > gpurun_out/ncu_x_trace.log 2>&1
ls -la gpurun_out/prof_wf_x_trace.ncu-rep
