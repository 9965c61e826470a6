"""Struct layouts: include/mrt_abi.h, the numpy views, and the oracle's restatement all agree with the layout
produced by compiling the reference's own renderer/Raytracing.h (tests/golden/abi_layout.json)."""
import json
import pathlib
import subprocess

import numpy as np

import mrt_b200._lib as L
import oracle as orc

ROOT = pathlib.Path(__file__).resolve().parents[1]
NAMES = {"Ray": L.RAY, "LightSamplingRay": L.LIGHT_RAY, "Intersection": L.INTERSECTION, "Vertex": L.VERTEX,
         "Material": L.MATERIAL, "TriangleReference": L.TRIANGLE_REFERENCE, "LightTriangle": L.LIGHT_TRIANGLE}


def test_numpy_views_match_reference_header(golden_dir):
    golden = json.loads((golden_dir / "abi_layout.json").read_text())["structs"]
    for name, dt in NAMES.items():
        g = golden[name]
        assert dt.itemsize == g["size"], name
        for field, off in g.items():
            if field in ("size", "align"):
                continue
            assert dt.fields[field][1] == off, (name, field)


def test_c_header_matches_reference_header(golden_dir, tmp_path):
    golden = json.loads((golden_dir / "abi_layout.json").read_text())
    cname = {"Spectrum": "mrt_spectrum", "SharedData": "mrt_shared_data", "Ray": "mrt_ray",
             "LightSamplingRay": "mrt_light_sampling_ray", "Intersection": "mrt_intersection", "Vertex": "mrt_vertex",
             "Material": "mrt_material", "TriangleReference": "mrt_triangle_reference", "LightTriangle": "mrt_light_triangle"}
    lines = ['#include "mrt.h"']
    for name, g in golden["structs"].items():
        lines.append(f'_Static_assert(sizeof({cname[name]}) == {g["size"]}, "{name} size");')
        lines.append(f'_Static_assert(_Alignof({cname[name]}) == {g["align"]}, "{name} align");')
        for field, off in g.items():
            if field not in ("size", "align"):
                lines.append(f'_Static_assert(offsetof({cname[name]}, {field}) == {off}, "{name}.{field}");')
    mac = golden["macros"]
    lines.append(f'_Static_assert(MRT_NOISE_DIMENSIONS == {mac["NOISE_DIMENSIONS"]} && MRT_MAX_PATH_LENGTH == {mac["MAX_PATH_LENGTH"]} '
                 f'&& MRT_SPECTRUM_SAMPLES == {mac["SPECTRUM_SAMPLES"]}, "macros");')
    lines.append(f'_Static_assert(MRT_MATERIAL_MIRROR == {mac["MATERIAL_MIRROR"]} && MRT_MATERIAL_SMOOTH_PLASTIC == {mac["MATERIAL_SMOOTH_PLASTIC"]} '
                 f'&& MRT_MATERIAL_SMOOTH_DIELECTRIC == {mac["MATERIAL_SMOOTH_DIELECTRIC"]} && MRT_MATERIAL_DIFFUSE == 0, "enum");')
    lines.append("int main(void) { return 0; }")
    src = tmp_path / "abi_check.c"
    src.write_text("\n".join(lines))
    subprocess.check_call(["gcc", "-std=c11", "-fsyntax-only", "-I", str(ROOT / "include"), str(src)])
    # epsilons / PI as float32
    assert np.float32(mac["DISTANCE_EPSILON"]) == np.float32(0.0001)
    assert np.float32(mac["ANGLE_EPSILON"]) == np.float32(0.00003807693583)
    assert np.float32(mac["PI"]) == np.float32(3.1415926)


def test_oracle_structs_match_reference_header(golden_dir):
    golden = json.loads((golden_dir / "abi_layout.json").read_text())["structs"]
    order = ["Ray", "LightSamplingRay", "Intersection", "Vertex", "Material", "TriangleReference", "LightTriangle",
             "SharedData", "Spectrum"]
    for i, name in enumerate(order):
        assert orc.lib().orc_sizeof(i) == golden[name]["size"], name
    for name, dt in {"Ray": orc.RAY, "LightSamplingRay": orc.LIGHT_RAY, "Intersection": orc.INTERSECTION,
                     "LightTriangle": orc.LIGHT_TRIANGLE}.items():
        for field, off in golden[name].items():
            if field not in ("size", "align"):
                assert dt.fields[field][1] == off, (name, field)


def test_library_exports_every_declared_symbol():
    import re
    header = (ROOT / "include" / "mrt.h").read_text()
    declared = set(re.findall(r"\b(mrt_[a-z_0-9]+)\s*\(", header))
    declared -= {"mrt_config", "mrt_ctx"}
    handle = L.lib()
    missing = [n for n in sorted(declared) if not hasattr(handle, n)]
    assert not missing, missing
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    assert handle.mrt_version() == 104
