#!/bin/bash
# round-2 GPU pass R (N GPUs, N = $1, default 2): the peer-memory reduction over real NVLink.
#   C driver: samples sharding with --reduce nccl and --reduce peer, twice each (peer must be bit-identical run to run), against one GPU
#   bench.py under torchrun: default (NCCL) with the peer reduction measured beside it through CUDA IPC, then --reduce peer
N=${1:-2}
DEV=$(seq -s, 0 $((N-1)))
mkdir -p gpurun_out
R=./metal-renderer_b200/bin/mrt_render
S="--scene spheres-1m --size 1920x1080 --seed 3 --frames 128"
{
timeout 300 $R $S --out gpurun_out/r_1gpu.pfm | tail -1
for k in 1 2; do
  timeout 300 $R $S --devices $DEV --shard samples --reduce nccl --out gpurun_out/r_nccl_$k.pfm | tail -1
  timeout 300 $R $S --devices $DEV --shard samples --reduce peer --out gpurun_out/r_peer_$k.pfm | tail -1
done
python scripts/compare_pfm.py gpurun_out/r_peer_1.pfm gpurun_out/r_peer_2.pfm
python scripts/compare_pfm.py gpurun_out/r_nccl_1.pfm gpurun_out/r_nccl_2.pfm
python scripts/compare_pfm.py gpurun_out/r_peer_1.pfm gpurun_out/r_nccl_1.pfm
python scripts/compare_pfm.py gpurun_out/r_peer_1.pfm gpurun_out/r_1gpu.pfm
} 2>&1 | tee gpurun_out/r_c_driver_${N}gpu.log
rm -f gpurun_out/r_*.pfm
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline \
  > gpurun_out/r_bench_${N}gpu_nccl.json 2> gpurun_out/r_bench_${N}gpu_nccl.err; tail -c 900 gpurun_out/r_bench_${N}gpu_nccl.json; tail -3 gpurun_out/r_bench_${N}gpu_nccl.err
[ "$N" -le 2 ] && timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 5 --warmup 3 --reduce peer --no-cpu-baseline \
  > gpurun_out/r_bench_${N}gpu_peer.json 2> gpurun_out/r_bench_${N}gpu_peer.err; tail -c 900 gpurun_out/r_bench_${N}gpu_peer.json; tail -3 gpurun_out/r_bench_${N}gpu_peer.err
python - <<PY
import json
for k in ("nccl", "peer"):
    try:
        d = json.loads(open("gpurun_out/r_bench_${N}gpu_%s.json" % k).read().strip().splitlines()[-1])
        print(k, "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "strong", {a: d["strong"].get(a) for a in ("ms_total", "allreduce_ms", "reduce", "other_reduce")}, "parity", d["parity"]["pass"] if d.get("parity") else None)
    except Exception as e:
        print(k, "no line:", e)
PY
