#!/bin/bash
# round-2 GPU pass Q (one GPU): the peer-memory reduction with several contexts on one device, the sharded step through it.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer_reduce.py tests/test_gpu_sharded.py tests/test_optics_extension.py -m gpu -q 2>&1 | tail -12 | tee gpurun_out/q_tests.log
