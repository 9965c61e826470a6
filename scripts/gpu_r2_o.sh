#!/bin/bash
# round-2 GPU pass O: the tests added since pass N (OBJ loader fixture, NaN-robust statistical tests of the extension) and a
# few images to look at: the same scenes with the reference's straight-through dielectric and with the optics extension.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_optics_extension.py tests/test_obj_loader.py -m gpu -q 2>&1 | tail -8 | tee gpurun_out/o_tests.log
R=metal-renderer_b200/bin/mrt_render
for s in own_box spheres-64k; do
  scene=$s; [ $s = own_box ] && scene=tests/golden/own_box.obj
  timeout 300 $R --scene $scene --size 640x360 --frames 512 --tone --srgb --out gpurun_out/o_${s}_reference.pfm | tail -1
  timeout 300 $R --scene $scene --size 640x360 --frames 512 --tone --srgb --optics all --spread 0.04 --out gpurun_out/o_${s}_optics.pfm | tail -1
  for k in reference optics; do python scripts/pfm_to_png.py gpurun_out/o_${s}_$k.pfm gpurun_out/o_${s}_$k.png && rm gpurun_out/o_${s}_$k.pfm; done
done
ls -la gpurun_out/*.png
