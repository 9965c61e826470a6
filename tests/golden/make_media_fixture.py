"""Build tests/golden/media_water.npz from the reference's own scene files (needs /root/reference; run here, not on the GPU box).

BASELINE.json configs[1] names `renderer/Media` scenes; the one the reference's MTL convention is exercised on is
Media/CornellBox-Water-plastic.obj (+ .mtl): Ks = (roughness, metalness, +-ior) decoded at renderer/Renderer.mm:278-329
into 2 smooth plastics (leftSphere ior 1.5 and water ior 1.33333, both from a NEGATIVE Ks.z), 1 mirror (rightSphere) and
6 diffuse materials of which `light` emits Ka = (10, 10, 10).  The fixture stores
  * the plastic scene as the ABI arrays OUR loader (csrc/obj_loader.cpp) produces — vertices, indices, triangle
    references, materials, light triangles with pdf / cdf / sentinel — so the GPU box can render the reference's real
    media without the reference tree, and
  * the material tables of the two sibling files (CornellBox-Water.mtl, -mirror.mtl) with their names, for the decode test.
The OBJ's 7,088 faces are triangles (CRLF line ends); SceneKit's own triangulation / vertex order is closed source, so
triangle ids are defined by our loader for the oracle and the CUDA path alike (oracle/README: parity unpinned there).
"""
import pathlib
import sys

import numpy as np

root = pathlib.Path(__file__).resolve().parents[2]
sys.path[:0] = [str(root / "metal-renderer_b200")]
import mrt_b200 as m  # noqa: E402

MEDIA = pathlib.Path("/root/reference/renderer/Media")


def material_names(mtl):
    return [line.split()[1] for line in open(mtl) if line.strip().startswith("newmtl")]


def first_use_order(obj):
    seen = []
    for line in open(obj, newline=None):
        p = line.split()
        if p and p[0] == "usemtl" and p[1] not in seen:
            seen.append(p[1])
    return seen


out = {}
for name in ("CornellBox-Water-plastic", "CornellBox-Water", "CornellBox-Water-mirror"):
    s = m.Scene.load_obj(str(MEDIA / f"{name}.obj"))
    out[f"{name}/materials"] = s.materials
    out[f"{name}/material_names"] = np.array(first_use_order(MEDIA / f"{name}.obj"))       # material index = order of first usemtl
    out[f"{name}/counts"] = np.array([len(s.vertices), s.triangle_count, s.light_count], np.uint32)
    if name == "CornellBox-Water-plastic":
        for field in ("vertices", "indices", "references", "lights"):
            out[f"{name}/{field}"] = getattr(s, field)
    print(name, "vertices", len(s.vertices), "triangles", s.triangle_count, "lights", s.light_count,
          "types", s.materials["materialType"].tolist(), "names", out[f"{name}/material_names"].tolist())
np.savez_compressed(root / "tests/golden/media_water.npz", **out)
print("wrote tests/golden/media_water.npz", (root / "tests/golden/media_water.npz").stat().st_size, "bytes")
