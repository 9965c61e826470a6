#!/bin/bash
# round-2 GPU pass W (N GPUs): the push-protocol peer reduction with a VMM-mapped exchange block in the C driver (no
# cudaDeviceEnablePeerAccess), against NCCL; then bench.py under torchrun (CUDA IPC block) with both reductions on record.
N=${1:-2}; DEV=$(seq -s, 0 $((N-1)))
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128"
{ timeout 300 $R $S --out gpurun_out/w_1gpu.pfm | tail -1 | cut -c150-420
for k in 1 2 3; do for red in nccl peer; do timeout 300 $R $S --devices $DEV --shard samples --reduce $red --out gpurun_out/w_${red}_$k.pfm | tail -1 | cut -c150-; done; done
python scripts/compare_pfm.py gpurun_out/w_peer_1.pfm gpurun_out/w_peer_2.pfm
python scripts/compare_pfm.py gpurun_out/w_peer_1.pfm gpurun_out/w_nccl_1.pfm
python scripts/compare_pfm.py gpurun_out/w_peer_1.pfm gpurun_out/w_1gpu.pfm; } 2>&1 | tee gpurun_out/w_c_driver_${N}gpu.log
rm -f gpurun_out/w_*.pfm
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline \
  > gpurun_out/w_bench_${N}gpu.json 2> gpurun_out/w_bench_${N}gpu.err; tail -3 gpurun_out/w_bench_${N}gpu.err
python - <<PY
import json
d = json.loads(open("gpurun_out/w_bench_${N}gpu.json").read().strip().splitlines()[-1])
print("value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "strong", {a: d["strong"].get(a) for a in ("ms_total", "allreduce_ms", "reduce", "other_reduce")}, "parity", d["parity"]["pass"])
PY
