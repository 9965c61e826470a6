"""Closed-form checks of every material branch THROUGH THE PIPELINE (intersect -> intersectionHandler), on a scene whose
answers can be written down by hand from renderer/Shaders.metal:105-212 and renderer/KernelHelpers.h:116-179:

  * two parallel mirror planes: the path zig-zags, direction = reflect(), throughput decays as prod(diffuse * cos theta)
    (the reference multiplies the mirror lobe by cos theta, KernelHelpers.h:170-174 — a kept quirk);
  * smooth dielectric below its Fresnel threshold: passes STRAIGHT through (no refraction, :156-158), bsdf = pdf = 1,
    ior := material ior, origin still offset along +n;
  * smooth plastic: coin flip `fresnel < noise.y` -> cosine lobe with throughput scale exactly `diffuse`, else mirror;
  * a path that runs into an emitter at bounce 0 adds emissive * throughput * 1 * mPdf (MIS weight 1: lPdf carries
    the previous-bounce-was-diffuse flag 0).

The expectations below are numpy restatements of the reference text, written independently of oracle/orc_shading.h and
csrc/device_shading.cuh (which are twin transcriptions by one author — bit-equality between them cannot catch a shared
misreading; these can).  The CPU test drives the oracle; the `gpu` twin drives the CUDA stage kernels through the C-ABI."""
import numpy as np
import pytest

import mrt_b200 as m

EPS = np.float32(0.0001)             # DISTANCE_EPSILON, Raytracing.h:16
W = H = 4
KD = np.float32([0.9, 0.8, 0.7])


def quad(vs, verts, idx, refs, mat, n):
    base = len(verts)
    for v in vs:
        verts.append((v, n))
    for a, b, c in ((0, 1, 2), (0, 2, 3)):
        idx.extend([base + a, base + b, base + c]); refs.append(((base + a, base + b, base + c), mat, m.NOT_A_LIGHT))


def two_planes(kind, ior=1.5, emissive_floor=False):
    """Floor y = 0 (normal +y) and ceiling y = 2 (normal -y) of material `kind`, 100 x 100 units, plus a small light far away."""
    verts, idx, refs = [], [], []
    quad([(-50, 0, -50), (-50, 0, 50), (50, 0, 50), (50, 0, -50)], verts, idx, refs, 0, (0, 1, 0))
    quad([(-50, 2, -50), (50, 2, -50), (50, 2, 50), (-50, 2, 50)], verts, idx, refs, 0, (0, -1, 0))
    base = len(verts)
    for v in ((40, 1, 40), (41, 1, 40), (40, 1, 41)):
        verts.append((v, (0, -1, 0)))
    idx.extend([base, base + 1, base + 2]); refs.append(((base, base + 1, base + 2), 1, 0))
    vertices = np.array(verts, dtype=m.VERTEX)
    references = np.array(refs, dtype=m.TRIANGLE_REFERENCE)
    materials = np.zeros(2, m.MATERIAL)
    materials[0]["diffuse"] = KD; materials[0]["ior"] = ior; materials[0]["materialType"] = kind
    if emissive_floor:
        materials[0]["emissive"] = (3, 2, 1)
    materials[1]["diffuse"] = 1; materials[1]["emissive"] = (5, 5, 5)
    lights = np.zeros(2, m.LIGHT_TRIANGLE)                     # one light + the sentinel (Renderer.mm:435-448)
    lights[0]["emissive"] = (5, 5, 5)
    for k, name in enumerate(("v1", "v2", "v3")):
        lights[0][name]["v"] = vertices["v"][base + k]; lights[0][name]["n"] = (0, -1, 0)
    lights[0]["area"], lights[0]["pdf"], lights[0]["cdf"], lights[0]["index"] = 0.5, 1.0, 0.0, 4
    lights[1]["cdf"] = 1.0
    if emissive_floor:                                         # the two floor triangles ARE the lights the path runs into
        lights = np.zeros(3, m.LIGHT_TRIANGLE)
        for t in range(2):
            lights[t]["emissive"] = (3, 2, 1)
            for k, name in enumerate(("v1", "v2", "v3")):
                lights[t][name] = vertices[references["tri"][t][k]]
            lights[t]["area"], lights[t]["pdf"], lights[t]["cdf"], lights[t]["index"] = 5000.0, 0.5, 0.5 * t, t
        lights[2]["cdf"] = 1.0
        references["lightTriangleIndex"][0:2] = (0, 1)
        references["lightTriangleIndex"][4] = m.NOT_A_LIGHT
        return m.Scene(vertices, np.array(idx, np.uint32), references, materials, lights, 2)
    return m.Scene(vertices, np.array(idx, np.uint32), references, materials, lights, 1)


def rays_all(direction, origin=(0.25, 1.0, 0.125)):
    r = np.zeros(W * H, m.RAY)
    r["origin"] = origin; r["direction"] = direction; r["maxDistance"] = np.inf
    r["throughput"] = 1; r["params"] = (1.0, 0.0, 0.0, 1.00029)            # rayGenerator, Shaders.metal:98-102
    return r


def fresnel_ref(cos_i, eta_i, eta_t):
    """KernelHelpers.h:7-21 in float64 (unpolarised dielectric Fresnel)."""
    sin_t2 = (eta_i / eta_t) ** 2 * (1 - cos_i ** 2)
    if sin_t2 > 1:
        return 1.0
    cos_t = np.sqrt(1 - sin_t2)
    rs = (eta_i * cos_i - eta_t * cos_t) / (eta_i * cos_i + eta_t * cos_t)
    rp = (eta_t * cos_i - eta_i * cos_t) / (eta_t * cos_i + eta_i * cos_t)
    return 0.5 * (rs * rs + rp * rp)


class OracleStages:
    def __init__(self, scene, depth=8):
        import oracle as orc
        self.o = orc.Oracle(max_path_length=depth, threads=1)
        self.o.init_raytracing(scene); self.o.drawable_size_will_change(W, H)

    def noise(self, tile):
        for s in range(3):
            self.o.update_noise(s, tile)

    def set_rays(self, rays): self.o.write_buffer(0, rays)
    def bounce(self, frame, i): self.o.intersect(0); self.o.shade(frame, i)
    def rays(self): return self.o.read_buffer(0)
    def light_rays(self): return self.o.read_buffer(1)


class CudaStages:
    def __init__(self, scene, depth=8):
        self.r = m.Renderer(max_path_length=depth, pipeline=m.PIPELINE_STAGED)
        self.r.init_raytracing(scene); self.r.drawable_size_will_change(W, H)

    def noise(self, tile):
        for s in range(3):
            self.r.update_noise(s, tile)

    def set_rays(self, rays): self.r.write_buffer(m.BUF_RAYS, rays)
    def bounce(self, frame, i): self.r.intersect(m.BUF_RAYS); self.r.shade(frame, i)
    def rays(self): return self.r.read_buffer(m.BUF_RAYS)
    def light_rays(self): return self.r.read_buffer(m.BUF_LIGHT_RAYS)


BACKENDS = [pytest.param(OracleStages, id="oracle"), pytest.param(CudaStages, id="cuda", marks=pytest.mark.gpu)]


def const_tile(x, y, z, w):
    return np.tile(np.float32([x, y, z, w]), 64 * 64)


@pytest.mark.parametrize("backend", BACKENDS)
def test_mirror_planes_throughput_decays_as_product_of_cosines(backend):
    b = backend(two_planes(m._lib.MATERIAL_MIRROR))
    b.noise(const_tile(0.3, 0.5, 0.25, 0.64))
    d = np.float32([0.6, -0.8, 0.0])
    b.set_rays(rays_all(d))
    want_dir, thr = d.astype(np.float64), np.ones(3)
    for k in range(1, 6):
        b.bounce(0, k - 1)
        n = np.array([0, 1, 0]) if k % 2 == 1 else np.array([0, -1, 0])
        want_dir = want_dir - 2 * np.dot(want_dir, n) * n               # reflect(i, n) = i - 2 dot(n, i) n
        thr = thr * KD.astype(np.float64) * 0.8                         # diffuse * bsdf / pdf = diffuse * cos / 1
        r = b.rays()
        assert np.allclose(r["direction"], want_dir, atol=1e-6), k
        assert np.allclose(r["throughput"], thr, rtol=2e-6), (k, r["throughput"][0], thr)
        assert (r["params"][:, 0] == 1).all() and (r["params"][:, 1] == 0).all() and (r["params"][:, 2] == k).all()
        y = float(EPS) if k % 2 == 1 else 2 - float(EPS)                # origin = hit + n * DISTANCE_EPSILON
        assert np.allclose(r["origin"][:, 1], y, atol=1e-6) and np.isinf(r["maxDistance"]).all()
        assert (r["radiance"] == 0).all()                               # the far light is never in the mirror direction
    assert thr[0] == pytest.approx((0.9 * 0.8) ** 5)


@pytest.mark.parametrize("backend", BACKENDS)
def test_dielectric_passes_straight_through_below_its_fresnel_threshold_and_reflects_above(backend):
    d = np.float32([0.6, -0.8, 0.0])
    f = fresnel_ref(0.8, 1.00029, 1.5)                                  # ray ior (raygen) -> material ior
    assert 0.04 < f < 0.05
    for noise_y, through in ((0.5, True), (0.01, False)):               # :151-166: `fI < noise.y` -> straight through
        b = backend(two_planes(m._lib.MATERIAL_SMOOTH_DIELECTRIC, ior=1.5))
        b.noise(const_tile(0.3, noise_y, 0.25, 0.64))
        b.set_rays(rays_all(d))
        b.bounce(0, 0)
        r = b.rays()
        if through:
            assert np.array_equal(r["direction"], np.tile(d, (W * H, 1)))            # no refraction in the reference
            assert np.allclose(r["throughput"], KD, rtol=1e-7) and (r["params"][:, 0] == 1).all()
            assert (r["params"][:, 3] == np.float32(1.5)).all()
        else:
            assert np.allclose(r["direction"], [0.6, 0.8, 0.0], atol=1e-6)
            assert np.allclose(r["throughput"], KD * np.float32(0.8), rtol=2e-6)
            assert (r["params"][:, 3] == np.float32(1.00029)).all()
        assert np.allclose(r["origin"][:, 1], float(EPS), atol=1e-7)                 # +n side either way
        assert (r["params"][:, 1] == 0).all() and (r["params"][:, 2] == 1).all()


@pytest.mark.parametrize("backend", BACKENDS)
def test_plastic_coin_flip_and_exact_diffuse_lobe_weight(backend):
    d = np.float32([0.6, -0.8, 0.0])
    for noise_y, diffuse_lobe in ((0.5, True), (0.01, False)):          # :143-149: diffuse lobe iff fresnel < noise.y
        b = backend(two_planes(m._lib.MATERIAL_SMOOTH_PLASTIC, ior=1.5))
        b.noise(const_tile(0.3, noise_y, 0.25, 0.64))
        b.set_rays(rays_all(d))
        b.bounce(0, 0)
        r = b.rays()
        if diffuse_lobe:
            assert np.array_equal(r["throughput"], np.tile(KD, (W * H, 1)))          # bsdf / pdf == 1 exactly
            cos_out = r["direction"][:, 1]
            assert np.allclose(cos_out, 0.8, atol=1e-6)                              # cosine lobe: cos = sqrt(noise.w) = 0.8
            assert np.allclose(r["params"][:, 0], cos_out / np.float32(3.1415926), rtol=1e-6)      # pdf = cos / PI
        else:
            assert np.allclose(r["direction"], [0.6, 0.8, 0.0], atol=1e-6)
            assert np.allclose(r["throughput"], KD * np.float32(0.8), rtol=2e-6) and (r["params"][:, 0] == 1).all()
        assert (r["params"][:, 1] == 0).all()                           # "previous bounce diffuse" is materialType == DIFFUSE only


@pytest.mark.parametrize("backend", BACKENDS)
def test_hitting_an_emitter_at_bounce_zero_adds_emission_with_weight_one(backend):
    b = backend(two_planes(m._lib.MATERIAL_DIFFUSE, emissive_floor=True))
    b.noise(const_tile(0.3, 0.5, 0.25, 0.64))
    b.set_rays(rays_all(np.float32([0.6, -0.8, 0.0])))
    b.bounce(0, 0)
    r = b.rays()
    # :186-201: mPdf = params.x = 1, lPdf = params.y (= 0 at bounce 0) * ..., weight = 1^2 / (1^2 + 0) = 1
    assert np.array_equal(r["radiance"], np.tile(np.float32([3, 2, 1]), (W * H, 1)))
    assert (r["params"][:, 1] == 1).all()                               # diffuse surface: the next hit gets MIS-weighted
    assert np.array_equal(r["throughput"], np.tile(KD, (W * H, 1)))


@pytest.mark.parametrize("backend", BACKENDS)
def test_light_sampling_ray_of_a_diffuse_hit(backend):
    b = backend(two_planes(m._lib.MATERIAL_DIFFUSE))
    b.noise(const_tile(0.3, 0.5, 0.25, 0.64))
    b.set_rays(rays_all(np.float32([0.6, -0.8, 0.0])))
    b.bounce(0, 0)
    lr = b.light_rays()                                                 # NEE ray towards the far light (Shaders.metal:145-170)
    assert (lr["targetIndex"] == 4).all() and np.isinf(lr["maxDistance"]).all()
    hit = np.array([0.25 + 0.75, 0.0, 0.125])
    assert np.allclose(lr["origin"], hit + [0, float(EPS), 0], atol=2e-5)        # interpolated from vertices at +-50
    to_light = lr["direction"].astype(np.float64)
    assert np.allclose(np.linalg.norm(to_light, axis=1), 1, atol=1e-6)
    # barycentric(noise.wx) picks a point of the light; every point of it lies within 1.5 of (40, 1, 40)
    want = np.array([40.3, 1, 40.3]) - hit
    assert (to_light @ (want / np.linalg.norm(want)) > 0.9995).all()
    # throughput = Le * kd * ray.throughput * weight * bsdf / lightPdf, all positive here (light faces down, hit faces up)
    assert (lr["throughput"] > 0).all()
