#!/bin/bash
# round-2 GPU pass N: the optics extension (8(f4)) — its own tests first, then the whole GPU suite (DeviceScene grew, so
# every kernel was recompiled), a bench line, and what the extension costs.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_optics_extension.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/n_ext_tests.log
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 | tee gpurun_out/n_gpu_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/n_bench.json 2> gpurun_out/n_bench.err; tail -c 600 gpurun_out/n_bench.json
{ timeout 600 python scripts/ext_cost.py spheres-1m 3 1920 1080 32; timeout 600 python scripts/ext_cost.py many-lights 5 1920 1080 32; } 2>&1 | tee gpurun_out/n_ext_cost.log
