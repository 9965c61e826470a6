"""Build tests/golden/mitsuba_reference.npz from the reference's own golden artefacts (needs /root/reference and cv2).

The reference ships Mitsuba renders of its scenes as the images its comparison mode is meant to be held against
(renderer/Media/reference/<scene>-<MAX_PATH_LENGTH>.exr, loaded at renderer/Renderer.mm:160-250; scene descriptions
renderer/Media/reference/*.xml: same camera as rayGenerator — eye (0,1,2.35), 90 degrees horizontal, 800x600).
They are the only result-pinning artefacts the reference has.  This script stores, for the diffuse scenes,
  * the EXR box-filtered 8x down to 100x75 (row 0 = bottom, RGB), with the pixels of Mitsuba's stamped label masked, and
  * the scene as the ABI arrays OUR loader produces from the reference's OBJ/MTL (so the test needs no reference tree).
The water scenes are left out: their plastic/dielectric models differ from Mitsuba's by construction.
"""
import os
import pathlib
import sys

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2  # noqa: E402
import numpy as np  # noqa: E402

root = pathlib.Path(__file__).resolve().parents[2]
sys.path[:0] = [str(root / "metal-renderer_b200")]
import mrt_b200 as m  # noqa: E402

MEDIA = pathlib.Path("/root/reference/renderer/Media")
F = 8
out = {}
for obj in ("cornellbox", "white-box"):
    s = m.Scene.load_obj(str(MEDIA / f"{obj}.obj"))
    for field in ("vertices", "indices", "references", "materials", "lights"):
        out[f"{obj}/{field}"] = getattr(s, field)
    out[f"{obj}/light_count"] = np.uint32(s.light_count)
for obj, depth in (("cornellbox", 2), ("cornellbox", 3), ("cornellbox", 8), ("white-box", 2)):
    img = cv2.imread(str(MEDIA / "reference" / f"{obj}-{depth}.exr"), cv2.IMREAD_UNCHANGED)
    assert img is not None and img.shape == (600, 800, 3)
    img = img[::-1, :, ::-1].astype(np.float64)                      # row 0 = bottom, BGR -> RGB
    label = (img >= 1000.0).any(-1)                                 # Mitsuba's text label is stamped at 1024
    h, w = label.shape
    total = np.where(label[..., None], 0.0, img).reshape(h // F, F, w // F, F, 3).sum((1, 3))
    count = (~label).reshape(h // F, F, w // F, F).sum((1, 3))
    out[f"{obj}-{depth}/image"] = (total / np.maximum(count, 1)[..., None]).astype(np.float32)
    out[f"{obj}-{depth}/valid"] = count == F * F
    print(obj, depth, "label pixels", int(label.sum()), "mean", float(out[f"{obj}-{depth}/image"][out[f"{obj}-{depth}/valid"]].mean()))
np.savez_compressed(root / "tests/golden/mitsuba_reference.npz", **out)
print("wrote tests/golden/mitsuba_reference.npz")
