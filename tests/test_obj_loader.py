"""csrc/obj_loader.cpp on an OBJ/MTL pair of our own (tests/golden/own_box.*, written by make_obj_fixture.py), so that the
loader — SURVEY.md 8(f1), the caller-side step in front of the path (Renderer.mm:257-454) — also runs where the
reference tree is absent (the GPU box).  CPU: every array the loader emits against an independent parser written from
the rules the loader documents; the Ks decode against the restatement of Renderer.mm:292-320 in test_media_fixture.
GPU: the loaded scene rendered bit-exactly against the oracle, with and without the optics extension (the file has a
closed glass block)."""
import pathlib

import numpy as np
import pytest

import mrt_b200 as m
from test_media_fixture import reference_decode

GOLDEN = pathlib.Path(__file__).resolve().parent / "golden"
OBJ = GOLDEN / "own_box.obj"


def parse_independently(path):
    """(positions+normals per vertex, triangles as vertex ids grouped by material, material names in first-use order)."""
    pos, nrm, verts, unique, names, groups = [], [], [], {}, [], {}
    cur = None
    for raw in path.read_bytes().split(b"\n"):
        tok = raw.decode().split()
        if not tok or tok[0].startswith("#"):
            continue
        if tok[0] == "v":
            pos.append(tuple(np.float32(x) for x in tok[1:4]))
        elif tok[0] == "vn":
            nrm.append(tuple(np.float32(x) for x in tok[1:4]))
        elif tok[0] == "usemtl":
            cur = tok[1]
            if cur not in groups:
                names.append(cur); groups[cur] = []
        elif tok[0] == "f":
            if cur is None:
                cur = "__default__"; names.append(cur); groups[cur] = []
            corners = []
            for c in tok[1:]:
                parts = c.split("/")
                vi = int(parts[0]); ni = int(parts[2]) if len(parts) > 2 and parts[2] else 0
                vi = vi if vi > 0 else len(pos) + vi + 1
                ni = ni if ni >= 0 else len(nrm) + ni + 1
                corners.append((vi, ni))
            face_n = None
            if any(ni == 0 for _, ni in corners):
                a, b, c = (np.array(pos[vi - 1], dtype=np.float32) for vi, _ in corners[:3])
                u, v = b - a, c - a
                n = np.array([u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]], dtype=np.float32)
                l = np.sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2], dtype=np.float32)
                face_n = tuple(n / l) if l > 0 else tuple(n)
            ids = []
            for vi, ni in corners:
                if ni == 0:
                    verts.append(pos[vi - 1] + face_n); ids.append(len(verts) - 1)
                else:
                    if (vi, ni) not in unique:
                        verts.append(pos[vi - 1] + nrm[ni - 1]); unique[(vi, ni)] = len(verts) - 1
                    ids.append(unique[(vi, ni)])
            for k in range(1, len(ids) - 1):
                groups[cur].append((ids[0], ids[k], ids[k + 1]))
    return np.array(verts, dtype=np.float32), groups, names


def parse_mtl(path):
    out, cur = {}, None
    for raw in path.read_bytes().split(b"\n"):
        tok = raw.decode().split()
        if not tok:
            continue
        if tok[0] == "newmtl":
            cur = tok[1]; out[cur] = {"Kd": (1, 1, 1), "Ka": (0, 0, 0), "Ks": (1, 0, 0)}
        elif tok[0] in ("Kd", "Ka", "Ks"):
            out[cur][tok[0]] = tuple(float(x) for x in tok[1:4])
    return out


def test_fixture_files_are_what_the_script_writes(tmp_path):
    import runpy, shutil
    before = (OBJ.read_bytes(), OBJ.with_suffix(".mtl").read_bytes())
    assert b"\r\n" in before[0] and b"\r\n" in before[1]
    shutil.copy(GOLDEN / "make_obj_fixture.py", tmp_path / "make_obj_fixture.py")
    runpy.run_path(str(tmp_path / "make_obj_fixture.py"))
    assert ((tmp_path / "own_box.obj").read_bytes(), (tmp_path / "own_box.mtl").read_bytes()) == before


def test_loader_output_equals_an_independent_parse():
    s = m.Scene.load_obj(str(OBJ))
    verts, groups, names = parse_independently(OBJ)
    assert names == ["__default__", "white", "red", "green", "light", "glass", "plastic", "mirror", "brushed", "frosted", "missing"]
    assert np.array_equal(s.vertices["v"], verts[:, :3]) and np.array_equal(s.vertices["n"], verts[:, 3:])
    tris = [t for n in names for t in groups[n]]
    owner = [i for i, n in enumerate(names) for _ in groups[n]]
    # 1 floor quad; ceiling + back wall + 1 later triangle; walls; light; 2 boxes; pentagon; 3 loose triangles
    assert [len(groups[n]) for n in names] == [2, 5, 2, 2, 2, 12, 12, 3, 1, 1, 1] and len(tris) == 43
    assert np.array_equal(s.indices.reshape(-1, 3), np.array(tris, dtype=np.uint32))
    assert np.array_equal(s.references["materialIndex"], np.array(owner, dtype=np.uint32))
    assert np.array_equal(s.references["tri"], np.array(tris, dtype=np.uint32))
    # materials: the reference's decode of Ks = (roughness, metalness, +-ior), Kd = diffuse, Ka = emission
    mtl = parse_mtl(OBJ.with_suffix(".mtl"))
    kinds = {}
    for mat, name in zip(s.materials, names):
        e = mtl.get(name, {"Kd": (1, 1, 1), "Ka": (0, 0, 0), "Ks": (1, 0, 0)})
        kind, ior = reference_decode(e["Ks"])
        assert (int(mat["materialType"]), mat["ior"]) == (kind, ior), name
        assert np.array_equal(mat["diffuse"], np.float32(e["Kd"])) and np.array_equal(mat["emissive"], np.float32(e["Ka"])), name
        kinds[name] = kind
    assert (kinds["glass"], kinds["plastic"], kinds["mirror"], kinds["brushed"], kinds["frosted"], kinds["red"]) == (3, 2, 1, 0, 0, 0)
    # light list (Renderer.mm:373-448): the two triangles of the `light` quad, area pdf, exclusive cdf, sentinel
    first = sum(len(groups[n]) for n in names[:4])
    assert s.light_count == 2 and s.lights["index"][:2].tolist() == [first, first + 1]
    assert np.allclose(s.lights["area"][:2], 0.18) and np.allclose(s.lights["pdf"][:2], 0.5)
    assert s.lights["cdf"].tolist() == [0.0, 0.5, 1.0] and s.lights["area"][2] == 0.0
    assert np.array_equal(s.lights["emissive"][0], np.float32([10, 9, 8]))
    assert s.references["lightTriangleIndex"][first: first + 2].tolist() == [0, 1]
    assert (np.delete(s.references["lightTriangleIndex"], [first, first + 1]) == m.NOT_A_LIGHT).all()


def test_loader_errors_are_reported(tmp_path):
    with pytest.raises(m.MrtError, match="cannot open"):
        m.Scene.load_obj(str(tmp_path / "absent.obj"))
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 7\n")
    with pytest.raises(m.MrtError, match="bad vertex index"):
        m.Scene.load_obj(str(bad))
    empty = tmp_path / "empty.obj"
    empty.write_text("v 0 0 0\n")
    with pytest.raises(m.MrtError, match="no faces"):
        m.Scene.load_obj(str(empty))


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [0, m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_ALIAS_LIGHTS])
def test_loaded_scene_renders_bit_exactly(flags):
    from helpers import bits_equal, setup_pair
    scene = m.Scene.load_obj(str(OBJ))
    r, o = setup_pair(scene, 160, 90, max_path_length=8)
    if flags:
        prob, alias = m.build_alias_table(scene.lights["pdf"][: scene.light_count])
        r.set_extensions(flags, 0.02); o.set_extensions(flags, 0.02, prob, alias)
    tiles = m.noise_tiles(6, 0, 6)
    r.draw(6, tiles); o.draw(6, tiles)
    img = r.save_current_image()
    assert bits_equal(img, o.save_current_image())
    seen = img[..., :3][np.isfinite(img[..., :3]).all(axis=2)]
    assert seen.mean() > 0.01                                            # the room is lit
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])
