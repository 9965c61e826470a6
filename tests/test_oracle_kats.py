"""Known-answer tests pinning the oracle (CPU restatement of Shaders.metal / KernelHelpers.h / Raytracing.h).

Two kinds of anchors:
  * tests/golden/ref_host_kats.json — outputs of the REFERENCE header's own host functions (compiled from
    /root/reference by oracle/ref_probe.cpp): balanceHeuristic, triangleSamplePDF must match bit-for-bit;
  * analytic identities for the MSL-only functions the reference cannot run here.
"""
import ctypes as C
import json
import math

import numpy as np
import pytest

import oracle as orc
from oracle import f32, _p

O = orc.lib()


def test_balance_heuristic_matches_reference_header(golden_dir):
    kats = json.loads((golden_dir / "ref_host_kats.json").read_text())["balanceHeuristic"]
    assert len(kats) == 100
    for f, g, want in kats:
        got = O.orc_balanceHeuristic(f, g)
        if isinstance(want, float) and math.isnan(want):
            assert math.isnan(got)
        else:
            assert np.float32(got).view(np.uint32) == np.float32(want).view(np.uint32), (f, g, got, want)
    assert O.orc_balanceHeuristic(1.0, 1.0) == 0.5


def test_triangle_sample_pdf_matches_reference_header(golden_dir):
    kats = json.loads((golden_dir / "ref_host_kats.json").read_text())["triangleSamplePDF"]
    for area, cos, dist, want in kats:
        got = O.orc_triangleSamplePDF(area, cos, dist)
        assert np.float32(got).view(np.uint32) == np.float32(want).view(np.uint32), (area, cos, dist)


def test_fresnel_known_answers():
    n, i = f32(0, 0, 1), f32(0, 0, 1)
    for eta in (1.33333, 1.5, 1.76):
        want = ((eta - 1) / (eta + 1)) ** 2                       # normal incidence
        assert O.orc_fresnel(_p(n), _p(i), 1.0, eta) == pytest.approx(want, rel=1e-6)
    # grazing incidence -> 1; total internal reflection (dense -> thin beyond the critical angle) -> exactly 1
    g = f32(math.sin(math.radians(89.999)), 0, math.cos(math.radians(89.999)))
    assert O.orc_fresnel(_p(n), _p(g), 1.0, 1.5) == pytest.approx(1.0, abs=1e-3)
    t = f32(math.sin(math.radians(60)), 0, math.cos(math.radians(60)))
    assert O.orc_fresnel(_p(n), _p(t), 1.5, 1.0) == 1.0
    # matched media -> 0
    assert O.orc_fresnel(_p(n), _p(t), 1.5, 1.5) == pytest.approx(0.0, abs=1e-12)


def test_barycentric_sums_to_one_and_matches_formula():
    rng = np.random.default_rng(0)
    out = np.zeros(3, np.float32)
    for x, y in rng.random((200, 2)).astype(np.float32):
        O.orc_barycentric(x, y, _p(out))
        assert abs(out.sum() - 1.0) < 3e-7 and (out >= 0).all()
        r1 = np.sqrt(np.float32(x))
        assert out[0] == np.float32(1) - r1 and out[2] == r1 * np.float32(y)


def test_orthonormal_basis():
    rng = np.random.default_rng(1)
    u, v = np.zeros(3, np.float32), np.zeros(3, np.float32)
    ns = rng.normal(size=(200, 3)); ns /= np.linalg.norm(ns, axis=1, keepdims=True)
    for n in list(ns.astype(np.float32)) + [f32(0, 0, 1), f32(0, 1, 0), f32(1, 0, 0), f32(0, 0.6, -0.8)]:
        O.orc_buildOrthonormalBasis(_p(n), _p(u), _p(v))
        for a, b in ((u, v), (u, n), (v, n)):
            assert abs(float(np.dot(a.astype(np.float64), b))) < 2e-6
        assert abs(np.linalg.norm(u) - 1) < 2e-6 and abs(np.linalg.norm(v) - 1) < 2e-6


def ulp_diff(a, b):
    ia, ib = int(np.float32(a).view(np.int32)), int(np.float32(b).view(np.int32))
    ia = ia if ia >= 0 else -(ia & 0x7FFFFFFF)
    ib = ib if ib >= 0 else -(ib & 0x7FFFFFFF)
    return abs(ia - ib)


def test_pinned_sincos_against_libm():
    s, c = C.c_float(), C.c_float()
    xs = np.concatenate([np.linspace(0, 2 * math.pi, 20001), np.random.default_rng(2).random(20000) * 2 * math.pi]).astype(np.float32)
    worst = 0.0
    for x in xs:
        O.orc_sincos(float(x), C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - math.sin(float(x))), abs(c.value - math.cos(float(x))))
    assert worst < 2.5e-7      # ~2 ulp at 1.0 in absolute terms
    O.orc_sincos(0.0, C.byref(s), C.byref(c))
    assert (s.value, c.value) == (0.0, 1.0)


def test_diffuse_bounce_is_cosine_hemisphere_sample():
    rng = np.random.default_rng(3)
    out = np.zeros(3, np.float32)
    for _ in range(300):
        n = rng.normal(size=3); n = (n / np.linalg.norm(n)).astype(np.float32)
        sx, sy = rng.random(2).astype(np.float32)
        O.orc_generateDiffuseBounce(sx, sy, _p(n), _p(out))
        assert abs(np.linalg.norm(out.astype(np.float64)) - 1) < 3e-6
        assert float(np.dot(out.astype(np.float64), n)) == pytest.approx(math.sqrt(sy), abs=3e-6)   # cos(theta) = sqrt(smp.y)


def test_reflect_and_interpolate():
    i, n, out = f32(0.6, -0.8, 0.0), f32(0, 1, 0), np.zeros(3, np.float32)
    O.orc_reflect(_p(i), _p(n), _p(out))
    assert np.allclose(out, [0.6, 0.8, 0.0], atol=1e-7)
    t = np.zeros(3, orc.VERTEX)
    t["v"] = [[1, 0, 0], [0, 1, 0], [0, 0, 1]]; t["n"] = [[0, 0, 2], [0, 0, 2], [0, 0, 2]]
    o = np.zeros(1, orc.VERTEX)
    O.orc_interpolate(_p(t[0:1]), _p(t[1:2]), _p(t[2:3]), 0.25, 0.5, _p(o))
    assert np.allclose(o["v"][0], [0.25, 0.5, 0.25]) and np.allclose(o["n"][0], [0, 0, 1])   # (u, v, 1-u-v), n normalised


def make_lights(areas):
    lights = np.zeros(len(areas) + 1, orc.LIGHT_TRIANGLE)
    total = np.float32(sum(np.float32(a) for a in areas)) if areas else np.float32(0)
    cdf = np.float32(0)
    for k, a in enumerate(areas):
        lights[k]["area"] = a; lights[k]["pdf"] = np.float32(a) / total; lights[k]["cdf"] = cdf
        cdf = np.float32(cdf + lights[k]["pdf"])
    lights[-1]["cdf"] = cdf; lights[-1]["pdf"] = 1.0
    return lights


def test_select_light_triangle_cdf_semantics():
    lights = make_lights([1.0, 1.0, 2.0])            # cdf 0, .25, .5 | sentinel 1.0
    pick = lambda xi: O.orc_selectLightTriangle(xi, _p(lights), 3)
    assert [pick(x) for x in (0.0, 0.2499, 0.25, 0.4999, 0.5, 0.999)] == [0, 0, 1, 1, 2, 2]
    assert pick(1.0) == 3                            # xi >= sum(pdf): the zero-area sentinel (KernelHelpers.h:52)
    assert O.orc_selectLightTriangle(0.3, _p(make_lights([])), 0) == 0


def test_sample_material_and_next_bounce_per_kind():
    out2, outd, bpi = np.zeros(2, np.float32), np.zeros(3, np.float32), np.zeros(3, np.float32)
    n, wI = f32(0, 1, 0), f32(0.6, -0.8, 0.0)
    wR = f32(0.6, 0.8, 0.0)                           # mirror direction
    noise = f32(0.3, 0.5, 0.25, 0.64)
    mat = np.zeros(1, orc.MATERIAL); mat["diffuse"] = 1; mat["ior"] = 1.5
    inv_pi = np.float32(1) / np.float32(3.1415926)
    # diffuse: (cos/pi, cos/pi)
    mat["materialType"] = 0
    O.orc_sampleMaterial(_p(mat), _p(wI), _p(wR), _p(n), _p(noise), _p(out2))
    assert out2[0] == out2[1] == inv_pi * np.float32(0.8)
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(noise), _p(outd), _p(bpi))
    assert bpi[0] == bpi[1] and bpi[2] == np.float32(1.00029) and outd[1] == pytest.approx(0.8, abs=1e-6)  # cos = sqrt(.64)
    # mirror: (cos, 1) only in the mirror direction, throughput scaled by cos (kept quirk)
    mat["materialType"] = 1
    O.orc_sampleMaterial(_p(mat), _p(wI), _p(wR), _p(n), _p(noise), _p(out2))
    assert tuple(out2) == (np.float32(0.8), 1.0)
    O.orc_sampleMaterial(_p(mat), _p(wI), _p(f32(0, 1, 0)), _p(n), _p(noise), _p(out2))
    assert tuple(out2) == (0.0, 1.0)
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(noise), _p(outd), _p(bpi))
    assert np.allclose(outd, wR, atol=1e-7) and bpi[1] == 1.0 and bpi[0] == pytest.approx(0.8, abs=1e-7)
    # dielectric: fresnel(…, 1.0, 1.5) at cos 0.8 ~ 0.0447 < noise.y = 0.5 -> straight through, (1,1), ior := material
    mat["materialType"] = 3
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(noise), _p(outd), _p(bpi))
    assert tuple(outd) == tuple(wI) and tuple(bpi) == (1.0, 1.0, 1.5)
    O.orc_sampleMaterial(_p(mat), _p(wI), _p(wR), _p(n), _p(noise), _p(out2))
    assert tuple(out2) == (0.0, 0.0)
    # ... and reflects when noise.y is below the Fresnel term
    low = f32(0.3, 0.01, 0.25, 0.64)
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(low), _p(outd), _p(bpi))
    assert np.allclose(outd, wR, atol=1e-7) and bpi[1] == 1.0
    # plastic: diffuse lobe when fresnel < noise.y, mirror otherwise
    mat["materialType"] = 2
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(noise), _p(outd), _p(bpi))
    assert bpi[0] == bpi[1]
    O.orc_generateNextBounce(_p(mat), _p(wI), 1.00029, _p(n), _p(low), _p(outd), _p(bpi))
    assert np.allclose(outd, wR, atol=1e-7) and bpi[1] == 1.0


def test_light_triangle_sample_pdf():
    sample = np.zeros(1, orc.VERTEX); sample["v"] = [0, 2, 0]; sample["n"] = [0, -1, 0]
    d = np.zeros(3, np.float32)
    pdf = O.orc_lightTriangleSamplePDF(0.5, 0.25, _p(f32(0, 0, 0)), _p(sample), _p(d))
    assert tuple(d) == (0.0, 1.0, 0.0) and pdf == pytest.approx(0.5 * 4.0 / (0.25 * 1.0))
    # back-facing or too close -> 0
    sample["n"] = [0, 1, 0]
    assert O.orc_lightTriangleSamplePDF(0.5, 0.25, _p(f32(0, 0, 0)), _p(sample), _p(d)) == 0.0
    sample["n"] = [0, -1, 0]
    assert O.orc_lightTriangleSamplePDF(0.5, 0.25, _p(f32(0, 2 - 5e-5, 0)), _p(sample), _p(d)) == 0.0


def test_triangle_test_convention_and_range():
    v0, v1, v2 = f32(1, 0, 0), f32(0, 1, 0), f32(0, 0, 0)
    tuv, hit = np.zeros(3, np.float32), C.c_uint32()
    # hit at P = 0.25*V0 + 0.5*V1 + 0.25*V2 -> coordinates (0.25, 0.5): weights of V0 and V1 (MPS convention)
    O.orc_triangle_test(_p(f32(0.25, 0.5, 1)), _p(f32(0, 0, -1)), 0.0, np.inf, _p(v0), _p(v1), _p(v2), _p(tuv), C.byref(hit))
    assert hit.value == 1 and tuple(tuv) == (1.0, 0.25, 0.5)
    # both faces hit (no culling)
    O.orc_triangle_test(_p(f32(0.25, 0.5, -1)), _p(f32(0, 0, 1)), 0.0, np.inf, _p(v0), _p(v1), _p(v2), _p(tuv), C.byref(hit))
    assert hit.value == 1
    # outside [minDistance, maxDistance], behind the origin, parallel, outside the triangle -> miss
    for o, d, tmin, tmax in ((f32(0.25, 0.5, 1), f32(0, 0, -1), 0.0, 0.5), (f32(0.25, 0.5, 1), f32(0, 0, -1), 1.5, 9.0),
                             (f32(0.25, 0.5, 1), f32(0, 0, 1), 0.0, np.inf), (f32(0.25, 0.5, 1), f32(1, 0, 0), 0.0, np.inf),
                             (f32(0.8, 0.8, 1), f32(0, 0, -1), 0.0, np.inf)):
        O.orc_triangle_test(_p(o), _p(d), tmin, tmax, _p(v0), _p(v1), _p(v2), _p(tuv), C.byref(hit))
        assert hit.value == 0
