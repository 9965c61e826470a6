#!/bin/bash
# round-2 GPU pass M: device time per frame for small batches (1, 2, 4 ... frames per mrt_render_frames call).
mkdir -p gpurun_out
{ timeout 600 python scripts/small_batches.py spheres-1m 3 1920 1080; timeout 600 python scripts/small_batches.py cornell-water 2 1920 1080; } 2>&1 | tee gpurun_out/m_small_batches.log
