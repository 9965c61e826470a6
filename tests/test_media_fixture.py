"""The reference's real media through the OBJ/MTL loader: Media/CornellBox-Water{-plastic,,-mirror}.obj.
CPU tests of the material decode against an independent restatement of renderer/Renderer.mm:278-329 applied to the MTL
values (copied below from the three .mtl files), of the light list (Renderer.mm:373-448) and of the committed fixture;
where /root/reference is present the loader is re-run and must reproduce the fixture bit for bit."""
import pathlib

import numpy as np
import pytest

import mrt_b200 as m

ROOT = pathlib.Path(__file__).resolve().parents[1]
FIXTURE = ROOT / "tests" / "golden" / "media_water.npz"
MEDIA = pathlib.Path("/root/reference/renderer/Media")
DIFFUSE, MIRROR, PLASTIC, DIELECTRIC = 0, 1, 2, 3

WALLS = {"floor": (0.7250, 0.7100, 0.6800), "ceiling": (0.7250, 0.7100, 0.6800), "backWall": (0.7250, 0.7100, 0.6800),
         "rightWall": (0.161, 0.133, 0.427), "leftWall": (0.6300, 0.0650, 0.0500), "light": (0.7250, 0.7100, 0.6800)}
# name -> (Kd, Ks, Ka) as written in the reference's .mtl files (Ks = roughness, metalness, +-ior; Ka = emission)
MTL = {
    "CornellBox-Water-plastic": {"leftSphere": ((1, 1, 1), (0.0, 0.0, -1.5), (0, 0, 0)), "rightSphere": ((1, 1, 1), (0.0, 1.0, 0.0), (0, 0, 0)),
                                 "water": ((1, 1, 1), (0.0, 0.0, -1.33333), (0, 0, 0))},
    "CornellBox-Water": {"leftSphere": ((1, 1, 1), (1.0, 0.0, 0.0), (0, 0, 0)), "rightSphere": ((1, 1, 1), (0.0, 0.0, -1.76), (0, 0, 0)),
                         "water": ((1, 1, 1), (0.0, 0.0, -1.33333), (0, 0, 0))},
    "CornellBox-Water-mirror": {"leftSphere": ((1, 1, 1), (1.0, 0.0, 0.0), (0, 0, 0)), "rightSphere": ((1, 1, 1), (1.0, 0.0, 0.0), (0, 0, 0)),
                                "water": ((1, 1, 1), (0.0, 1.0, 0.0), (0, 0, 0))},
}
for table in MTL.values():
    for wall, kd in WALLS.items():
        table[wall] = (kd, (1.0, 0.0, 0.0), (10, 10, 10) if wall == "light" else (0, 0, 0))


def reference_decode(ks):
    """Renderer.mm:292-320, restated independently of csrc/obj_loader.cpp: -> (materialType, ior)."""
    roughness, metalness, ior = (np.float32(x) for x in ks)
    kind = DIFFUSE                                              # Material{} is zero-initialised: type 0
    if metalness > 0:
        if roughness == 0:
            kind = MIRROR
    elif roughness == 1:
        kind = DIFFUSE
    elif ior <= 0:
        ior = abs(ior)
        kind = PLASTIC if roughness == 0 else DIFFUSE
    else:
        kind = DIELECTRIC if roughness == 0 else DIFFUSE
    return kind, np.float32(ior)


@pytest.mark.parametrize("name", sorted(MTL))
def test_material_decode_of_the_reference_media(name):
    with np.load(FIXTURE) as z:
        mats, names, counts = z[f"{name}/materials"], z[f"{name}/material_names"].tolist(), z[f"{name}/counts"].tolist()
    assert counts == [3666, 7088, 2]                            # the OBJ's faces are triangles: 7,088, not 14,176
    assert len(mats) == 9 and sorted(names) == sorted(MTL[name])
    for mat, mname in zip(mats, names):
        kd, ks, ka = MTL[name][mname]
        kind, ior = reference_decode(ks)
        assert int(mat["materialType"]) == kind, mname
        assert mat["ior"] == ior, mname
        assert np.array_equal(mat["diffuse"], np.float32(kd)) and np.array_equal(mat["emissive"], np.float32(ka)), mname
    kinds = sorted(int(k) for k in mats["materialType"])
    if name == "CornellBox-Water-plastic":                      # 2 smooth plastic (negative Ks.z), 1 mirror, 6 diffuse
        assert kinds == [DIFFUSE] * 6 + [MIRROR] + [PLASTIC] * 2
        by = dict(zip(names, mats))
        assert by["leftSphere"]["ior"] == np.float32(1.5) and by["water"]["ior"] == np.float32(1.33333)
        assert np.array_equal(by["light"]["emissive"], np.float32([10, 10, 10]))


def test_plastic_scene_fixture_is_a_valid_scene_with_the_reference_light_list():
    s = m.Scene.from_npz(FIXTURE, "CornellBox-Water-plastic")
    assert (len(s.vertices), s.triangle_count, s.light_count) == (3666, 7088, 2)
    light_mat = int(np.nonzero(s.materials["emissive"].sum(axis=1) > 0)[0][0])
    emitters = np.nonzero(s.references["materialIndex"] == light_mat)[0]
    assert len(emitters) == 2 and sorted(s.lights["index"][:2].tolist()) == sorted(emitters.tolist())
    assert sorted(s.references["lightTriangleIndex"][emitters].tolist()) == [0, 1]
    assert (s.references["lightTriangleIndex"][np.setdiff1d(np.arange(s.triangle_count), emitters)] == m.NOT_A_LIGHT).all()
    # Renderer.mm:435-448: pdf = area / total area, exclusive cdf, sentinel with cdf = 1
    area = s.lights["area"][:2]
    assert np.allclose(s.lights["pdf"][:2], area / area.sum(), rtol=1e-6)
    assert s.lights["cdf"][0] == 0 and np.isclose(s.lights["cdf"][1], s.lights["pdf"][0]) and s.lights["cdf"][2] == 1
    assert (np.diff(s.lights["cdf"]) >= 0).all()


@pytest.mark.skipif(not MEDIA.exists(), reason="reference tree not present (GPU box)")
def test_loader_reproduces_the_fixture_from_the_reference_files():
    s = m.Scene.load_obj(str(MEDIA / "CornellBox-Water-plastic.obj"))
    f = m.Scene.from_npz(FIXTURE, "CornellBox-Water-plastic")
    for field in ("vertices", "indices", "references", "materials", "lights"):
        assert getattr(s, field).tobytes() == getattr(f, field).tobytes(), field
    for name in ("CornellBox-Water", "CornellBox-Water-mirror"):
        with np.load(FIXTURE) as z:
            assert m.Scene.load_obj(str(MEDIA / f"{name}.obj")).materials.tobytes() == z[f"{name}/materials"].tobytes()
