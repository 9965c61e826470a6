#!/bin/bash
# round-2 GPU pass J: the 10 M-triangle scene at 4K is latency-bound (L2 hit 60 %, long scoreboard 6.5), unlike the 1 M scene:
# sweep what trades issue efficiency for memory-level parallelism — blocks per SM, refill threshold, vote thresholds, the
# collapse's triangle cost / leaf size, the L2 persisting window — on the experiments build.  Trace ms from the per-launch events.
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render_exp
S="--scene spheres-10m --size 3840x2160 --seed 4 --frames 8"
run() { echo "== $*"; env "$@" MRT_WAVE_LOG=/dev/null timeout 300 $R $S --pipeline wavefront 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-12s mrays %8.1f device_ms %8.2f trace_ms %8.2f nodes8 %d depth %d'%(d['scene'],d['mrays_per_s'],d['device_ms'],d['profile_trace_ms'],d['bvh8_nodes'],d['bvh8_depth']))"; }
{
export MRT_BVH_BINS=64          # the tree the sampled selection picks for this scene (saves the second build per run)
run A=1
run A=2
for mb in 6 9 10 12; do run MRT_TRACE_MINBLOCKS=$mb; done
for rf in 16 20 28 32; do run MRT_TRACE_REFILL=$rf; done
run MRT_TRACE_LEAFMIN=6 MRT_TRACE_TRIRATIO=5
run MRT_TRACE_LEAFMIN=14 MRT_TRACE_TRIRATIO=9
run MRT_TRACE_LEAFMIN=32 MRT_TRACE_TRIRATIO=0
for cp in 0.7 1.5 2.0; do run MRT_BVH8_CPRIM=$cp; done
run MRT_BVH8_MAXLEAF=2
run MRT_BVH8_MAXLEAF=1
run MRT_L2_PERSIST=1
run MRT_L2_PERSIST=48
run MRT_TRACE_DEFER=1
run MRT_TRACE_TQ=32
run MRT_SHADE_MINBLOCKS=3
run A=3
} 2>&1 | tee gpurun_out/j_sweep_10m.log
