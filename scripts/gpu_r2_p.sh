#!/bin/bash
# round-2 GPU pass P: extension tests incl. the continuous-spectrum flag and the 1080p cases, OBJ fixture render, images.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_optics_extension.py tests/test_obj_loader.py tests/test_abi_layout.py -m gpu -q 2>&1 | tail -8 | tee gpurun_out/p_tests.log
R=metal-renderer_b200/bin/mrt_render
timeout 300 $R --scene spheres-64k --size 640x360 --frames 1024 --tone --srgb --optics spectrum --spread 0.04 --out gpurun_out/p.pfm | tail -1
python scripts/pfm_to_png.py gpurun_out/p.pfm gpurun_out/p_spheres-64k_spectrum.png && rm gpurun_out/p.pfm
timeout 300 $R --scene spheres-64k --size 640x360 --frames 1024 --tone --srgb --optics dispersion --spread 0.04 --out gpurun_out/p.pfm | tail -1
python scripts/pfm_to_png.py gpurun_out/p.pfm gpurun_out/p_spheres-64k_dispersion.png && rm gpurun_out/p.pfm
