"""Optics extension: Snell refraction, one-channel dispersion, alias-table light selection.

BEYOND THE REFERENCE: the reference continues straight through dielectrics (KernelHelpers.h:151-166) and scans the light
CDF (:49-54), so nothing here can be pinned against it.  The checks are
  * closed forms on the CPU restatement (oracle/orc_shading.h refractiveBounce, selectLightAlias): Snell's law, a slab
    returns the incoming direction, total internal reflection beyond the critical angle, radiance scale eta^2 and its
    round trip, the alias table's marginals equal the pdfs;
  * the CUDA path equals that restatement bit for bit, dispatch by dispatch (staged) and on images (wavefront);
  * flags = 0 is the reference's behaviour, bit for bit; the alias estimator agrees with the CDF estimator in the mean.
"""
import math

import numpy as np
import pytest

import mrt_b200 as m
import oracle as orc
from oracle import f32, _p

O = orc.lib()
AIR = np.float32(1.00029)


def refract(mat_ior, wI, n, noise, ray_ior=AIR, flags=m.EXT_REFRACTION, dispersion=0.0, throughput=(1, 1, 1)):
    mat = np.zeros(1, orc.MATERIAL); mat["diffuse"] = 1; mat["ior"] = mat_ior; mat["materialType"] = 3
    out, bpi, scale = f32(0, 0, 0), f32(0, 0, 0), f32(0, 0, 0)
    O.orc_refractiveBounce(_p(mat), flags, dispersion, _p(f32(*wI)), float(ray_ior), _p(f32(*throughput)), _p(f32(*n)),
                           _p(f32(*noise)), _p(out), _p(bpi), _p(scale))
    return out, bpi, scale


@pytest.mark.parametrize("deg", [0.0, 10.0, 35.0, 60.0, 85.0])
def test_snell_and_slab_round_trip(deg):
    th = math.radians(deg)
    wI = (math.sin(th), -math.cos(th), 0.0)                 # travelling down onto a surface with normal +y
    n = (0.0, 1.0, 0.0)
    noise = (0.5, 0.999, 0.1, 0.1)                          # noise.y above any Fresnel value short of 1: transmit
    d1, bpi1, s1 = refract(1.5, wI, n, noise)
    sin_t = math.sin(th) * float(AIR) / 1.5
    assert d1[0] == pytest.approx(sin_t, abs=2e-7) and d1[1] == pytest.approx(-math.sqrt(1 - sin_t ** 2), abs=2e-7) and d1[2] == 0.0
    assert bpi1[0] == pytest.approx((float(AIR) / 1.5) ** 2, rel=1e-6) and bpi1[1] == 1.0 and bpi1[2] == np.float32(1.5)
    assert list(s1) == [1.0, 1.0, 1.0]
    # the lower face of the slab has the outward normal -y: the path leaves into air, parallel to where it came from
    d2, bpi2, _ = refract(1.5, d1, (0.0, -1.0, 0.0), noise, ray_ior=bpi1[2])
    assert np.allclose(d2, f32(*wI), atol=3e-7)
    assert bpi2[2] == AIR and float(bpi1[0]) * float(bpi2[0]) == pytest.approx(1.0, rel=1e-6)     # radiance scale undone


def test_total_internal_reflection_and_fresnel_reflection():
    n = (0.0, 1.0, 0.0)                                     # outward normal; the ray comes from inside (travels up)
    crit = math.asin(float(AIR) / 1.5)
    for deg_off, reflects in [(-3.0, False), (3.0, True)]:
        th = crit + math.radians(deg_off)
        wI = (math.sin(th), math.cos(th), 0.0)
        d, bpi, _ = refract(1.5, wI, n, (0.5, 0.9999, 0, 0), ray_ior=1.5)
        if reflects:
            assert np.allclose(d, f32(wI[0], -wI[1], 0.0), atol=2e-7) and bpi[2] == np.float32(1.5)   # still inside
            assert bpi[0] == 1.0 and bpi[1] == 1.0                                                    # lossless
        else:
            assert d[1] > 0 and bpi[2] == AIR
            assert d[0] == pytest.approx(math.sin(th) * 1.5 / float(AIR), abs=3e-6)
    # Fresnel reflection from outside: noise.y = 0 is below any reflectance > 0 -> mirror about the normal, index unchanged
    wI = (0.6, -0.8, 0.0)
    d, bpi, _ = refract(1.5, wI, n, (0.5, 0.0, 0, 0))
    assert np.allclose(d, f32(0.6, 0.8, 0.0), atol=2e-7) and list(bpi) == [1.0, 1.0, AIR]


def test_dispersion_keeps_one_channel_and_bends_blue_more():
    wI, n = (0.6, -0.8, 0.0), (0.0, 1.0, 0.0)
    flags = m.EXT_REFRACTION | m.EXT_DISPERSION
    sines = []
    for c, x in enumerate((0.1, 0.5, 0.9)):                 # noise.x picks the channel among the 3 live ones
        d, bpi, scale = refract(1.5, wI, n, (x, 0.999, 0, 0), flags=flags, dispersion=0.02)
        want = [0.0, 0.0, 0.0]; want[c] = 3.0
        assert list(scale) == want
        assert bpi[2] == np.float32(np.float32(1.5) + np.float32(0.02) * np.float32(c - 1))
        assert d[0] == pytest.approx(0.6 * float(AIR) / float(bpi[2]), abs=3e-7)
        sines.append(float(d[0]))
    assert sines[0] > sines[1] > sines[2]                   # red is deviated least, blue most
    # two live channels: weight 2, never the dead one; one live channel: no choice, weight 1
    for x in (0.1, 0.9):
        _, _, scale = refract(1.5, wI, n, (x, 0.999, 0, 0), flags=flags, dispersion=0.02, throughput=(0.5, 0.0, 0.25))
        assert list(scale) == ([2.0, 0.0, 0.0] if x < 0.5 else [0.0, 0.0, 2.0])
    _, bpi, scale = refract(1.5, wI, n, (0.7, 0.999, 0, 0), flags=flags, dispersion=0.02, throughput=(0.0, 0.3, 0.0))
    assert list(scale) == [0.0, 1.0, 0.0] and bpi[2] == np.float32(1.5)
    # leaving: no new choice, the carried index decides
    d, bpi, scale = refract(1.5, (0.3, 0.95393920141, 0.0), n, (0.1, 0.999, 0, 0), ray_ior=1.52, flags=flags, dispersion=0.02,
                            throughput=(0.0, 0.0, 1.0))
    assert list(scale) == [1.0, 1.0, 1.0] and bpi[2] == AIR and d[0] == pytest.approx(0.3 * 1.52 / float(AIR), abs=3e-6)
    # continuous spectrum: the index moves inside the band with noise.w; the band centre (w = 0.5) is the discrete value
    cont = flags | m.EXT_CONTINUOUS_SPECTRUM
    for w, off in ((0.5, 0.0), (0.0, -0.5), (0.75, 0.25)):
        _, bpi, scale = refract(1.5, wI, n, (0.9, 0.999, 0, w), flags=cont, dispersion=0.02)      # noise.x = 0.9: blue band
        assert list(scale) == [0.0, 0.0, 3.0]
        assert bpi[2] == np.float32(np.float32(1.5) + np.float32(0.02) * np.float32(np.float32(1.0) + (np.float32(w) - np.float32(0.5))))
    # without the DISPERSION bit the coefficient is ignored
    _, bpi, scale = refract(1.5, wI, n, (0.1, 0.999, 0, 0), flags=m.EXT_REFRACTION, dispersion=0.02)
    assert list(scale) == [1.0, 1.0, 1.0] and bpi[2] == np.float32(1.5)


def alias_marginals(prob, alias):
    n = len(prob)
    p = prob.astype(np.float64).copy()
    np.add.at(p, alias, 1.0 - prob.astype(np.float64))
    return p / n


@pytest.mark.parametrize("n,seed", [(1, 0), (2, 1), (7, 2), (1000, 3), (2048, 4)])
def test_alias_table_marginals_equal_the_pdfs(n, seed):
    rng = np.random.default_rng(seed)
    pdf = rng.uniform(0.0, 1.0, n).astype(np.float32) ** 3
    if n >= 7:
        pdf[rng.integers(0, n, n // 5)] = 0.0               # zero-area emitters are never selected
    pdf /= pdf.sum(dtype=np.float64)
    prob, alias = m.build_alias_table(pdf)
    assert alias.max() < n and prob.min() >= 0.0 and prob.max() <= 1.0
    assert np.allclose(alias_marginals(prob, alias), pdf, atol=2e-7)
    assert np.all(prob[pdf == 0] == 0.0) and not np.any(np.isin(alias[prob < 1], np.flatnonzero(pdf == 0)))
    # the selection function over a fine grid of xi reproduces them; xi = 1 stays in range
    xs = (np.arange(200 * n) + 0.5) / (200 * n)
    picks = np.array([O.orc_selectLightAlias(float(x), _p(prob), _p(alias), n) for x in xs[:: max(1, n // 50)]])
    assert picks.max() < n and not np.any(pdf[picks] == 0)
    if n <= 7:
        freq = np.bincount([O.orc_selectLightAlias(float(x), _p(prob), _p(alias), n) for x in xs], minlength=n) / len(xs)
        assert np.allclose(freq, pdf, atol=2.0 / 200)
    assert O.orc_selectLightAlias(1.0, _p(prob), _p(alias), n) < n and O.orc_selectLightAlias(0.0, _p(prob), _p(alias), n) < n


def test_scene_light_pdfs_survive_the_table():
    scene = m.Scene.procedural("many-lights", 0)
    pdf = scene.lights["pdf"][: scene.light_count]
    prob, alias = m.build_alias_table(pdf)
    assert np.allclose(alias_marginals(prob, alias), pdf / pdf.sum(dtype=np.float64), atol=1e-7)


def water_as_glass(name="cornell-water"):
    """The procedural scene with its plastic water surface (ior 1.33333) turned into a smooth dielectric."""
    scene = m.Scene.procedural(name, 0)
    mats = scene.materials
    mats["materialType"][np.isclose(mats["ior"], 1.33333)] = 3
    assert (mats["materialType"] == 3).sum() == 1
    return scene


def test_oracle_with_no_flags_is_the_reference_path():
    scene = water_as_glass()
    tiles = m.noise_tiles(5, 0, 2)
    imgs = []
    for flags in (None, 0):
        o = orc.Oracle(max_path_length=5)
        o.init_raytracing(scene); o.drawable_size_will_change(48, 32)
        if flags is not None:
            o.set_extensions(flags)
        o.draw(2, tiles)
        imgs.append(o.save_current_image())
    assert np.array_equal(imgs[0].view(np.uint32), imgs[1].view(np.uint32))
    o = orc.Oracle(max_path_length=5)
    o.init_raytracing(scene); o.drawable_size_will_change(48, 32)
    o.set_extensions(m.EXT_REFRACTION)
    o.draw(2, tiles)
    assert not np.array_equal(o.save_current_image().view(np.uint32), imgs[0].view(np.uint32))


# ---- CUDA path against the restatement ---------------------------------------------------------------------------
from helpers import bits_equal, mismatch_report, setup_pair  # noqa: E402

ALL = m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_ALIAS_LIGHTS | m.EXT_CONTINUOUS_SPECTRUM
COMBOS = [pytest.param(m.EXT_REFRACTION, 0.0, id="refraction"), pytest.param(m.EXT_REFRACTION | m.EXT_DISPERSION, 0.03, id="dispersion"),
          pytest.param(m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_CONTINUOUS_SPECTRUM, 0.03, id="spectrum"),
          pytest.param(m.EXT_ALIAS_LIGHTS, 0.0, id="alias"), pytest.param(ALL, 0.02, id="all")]


def assert_bits(got, want, what):
    assert bits_equal(got, want), f"{what}: {mismatch_report(got, want) if got.dtype.fields else 'bits differ'}"


def enable(r, o, scene, flags, dispersion):
    prob, alias = m.build_alias_table(scene.lights["pdf"][: scene.light_count])
    r.set_extensions(flags, dispersion)
    o.set_extensions(flags, dispersion, prob, alias)


@pytest.mark.gpu
@pytest.mark.parametrize("flags,dispersion", COMBOS)
def test_extension_every_stage_buffer_matches_restatement(flags, dispersion):
    scene = water_as_glass()
    w, h, L = 96, 64, 6
    r, o = setup_pair(scene, w, h, max_path_length=L)
    enable(r, o, scene, flags, dispersion)
    tiles = m.noise_tiles(33, 0, 2)
    for f in range(2):
        r.update_noise(f % 3, tiles[f]); o.update_noise(f % 3, tiles[f])
        r.raygen(f); o.raygen(f)
        for i in range(L):
            r.intersect(m.BUF_RAYS); o.intersect(0)
            assert_bits(r.read_buffer(m.BUF_INTERSECTIONS), o.read_buffer(2), f"f{f} i{i} primary hits")
            r.shade(f, i); o.shade(f, i)
            assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"f{f} i{i} shade rays")
            assert_bits(r.read_buffer(m.BUF_LIGHT_RAYS), o.read_buffer(1), f"f{f} i{i} light rays")
            r.intersect(m.BUF_LIGHT_RAYS); o.intersect(1)
            r.shadow_resolve(); o.shadow_resolve()
            assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"f{f} i{i} resolve")
        r.accumulate(f); o.accumulate(f)
        assert_bits(r.read_buffer(m.BUF_IMAGE), o.read_buffer(3), f"f{f} image")


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h,frames", [("cornell-water", 128, 72, 5), ("spheres-64k", 128, 72, 4), ("many-lights", 96, 54, 4)])
@pytest.mark.parametrize("flags,dispersion", COMBOS)
def test_extension_wavefront_image_matches_restatement(name, w, h, frames, flags, dispersion):
    """Closed glass spheres (spheres-64k, many-lights: enter, total internal reflection, leave) and the open water surface."""
    scene = water_as_glass() if name == "cornell-water" else m.Scene.procedural(name, 5)
    r, o = setup_pair(scene, w, h, max_path_length=8, renderer_kwargs=dict(pipeline=m.PIPELINE_WAVEFRONT))
    enable(r, o, scene, flags, dispersion)
    tiles = m.noise_tiles(9, 0, frames)
    r.draw(frames, tiles); o.draw(frames, tiles)
    assert_bits(r.save_current_image(), o.save_current_image(), name)
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])


@pytest.mark.gpu
def test_flags_zero_restores_the_reference_bits_and_bad_flags_are_rejected():
    scene = m.Scene.procedural("spheres-64k", 5)
    w, h, frames = 96, 54, 3
    tiles = m.noise_tiles(4, 0, frames)
    r, o = setup_pair(scene, w, h)
    o.draw(frames, tiles)
    want = o.save_current_image()
    r.set_extensions(ALL, 0.02)
    r.draw(frames, tiles)
    assert not bits_equal(r.save_current_image(), want)                 # glass spheres refract now
    r.set_extensions(0)
    r.drawable_size_will_change(w, h)                                   # restart the accumulation
    init = m.noise_tile(1, 0xFFFFFFFF)
    for s in range(3):
        r.update_noise(s, init)
    r.draw(frames, tiles)
    assert_bits(r.save_current_image(), want, "after set_extensions(0)")
    with pytest.raises(m.MrtError, match="unknown extension bit"):
        r.set_extensions(16)
    with pytest.raises(m.MrtError, match="needs MRT_EXT_DISPERSION"):
        r.set_extensions(m.EXT_REFRACTION | m.EXT_CONTINUOUS_SPECTRUM)
    with pytest.raises(m.MrtError, match="needs MRT_EXT_REFRACTION"):
        r.set_extensions(m.EXT_DISPERSION, 0.01)
    with pytest.raises(m.MrtError, match="dispersion"):
        r.set_extensions(m.EXT_REFRACTION | m.EXT_DISPERSION, 0.75)
    r.set_extensions(m.EXT_REFRACTION)
    r.set_pipeline(m.PIPELINE_MEGAKERNEL)
    with pytest.raises(m.MrtError, match="wavefront and staged"):
        r.draw(1, tiles[:1])


@pytest.mark.gpu
def test_alias_estimator_agrees_with_the_cdf_estimator_in_the_mean():
    """Same integrand, different light samples: the two images agree within Monte-Carlo noise (256 spp, 2,048 emitters)."""
    scene = m.Scene.procedural("many-lights", 0)
    w, h, frames = 64, 36, 256
    tiles = m.noise_tiles(12, 0, frames)
    imgs = []
    for flags in (0, m.EXT_ALIAS_LIGHTS):
        r = m.Renderer(max_path_length=4)
        r.init_raytracing(scene); r.drawable_size_will_change(w, h)
        for slot in range(3):                                    # the reference fills its three slots at start-up (Renderer.mm:115-126)
            r.update_noise(slot, m.noise_tile(12, 0xFFFFFFFF))
        r.set_extensions(flags)
        r.draw(frames, tiles)
        imgs.append(r.save_current_image()[..., :3].astype(np.float64))
    # a pixel whose running mean met a 0/0 sample (noise.w == 0, Shaders.metal:209 — kept quirk) stays NaN: leave those out
    ok = np.isfinite(imgs[0]).all(axis=2) & np.isfinite(imgs[1]).all(axis=2)
    assert ok.mean() > 0.99
    means = [img[ok].mean(axis=0) for img in imgs]
    assert np.allclose(means[0], means[1], rtol=0.02), means


@pytest.mark.gpu
def test_dispersion_does_not_change_the_mean_image():
    """Keeping one channel with weight = number of live channels is unbiased up to the per-channel index: a spread of
    +-0.03 around 1.5 moves the mean of an image full of glass spheres by Monte-Carlo noise only."""
    scene = m.Scene.procedural("spheres-64k", 2)
    r = m.Renderer(max_path_length=8)
    r.init_raytracing(scene); r.drawable_size_will_change(64, 36)
    tiles = m.noise_tiles(2, 0, 64)
    init = m.noise_tile(2, 0xFFFFFFFF)
    for slot in range(3):
        r.update_noise(slot, init)
    r.set_extensions(m.EXT_REFRACTION)
    r.draw(64, tiles)
    a = r.save_current_image()[..., :3]
    r.set_extensions(m.EXT_REFRACTION | m.EXT_DISPERSION, 0.03)
    r.drawable_size_will_change(64, 36)
    for slot in range(3):
        r.update_noise(slot, init)
    r.draw(64, tiles)
    b = r.save_current_image()[..., :3]
    ok = np.isfinite(a).all(axis=2) & np.isfinite(b).all(axis=2)        # 0/0 samples of the reference path stay NaN
    assert ok.mean() > 0.99                 # (small negative values are the reference's: its NEE cosine is not clamped, KernelHelpers.h:61)
    # channel selection is unbiased: the image mean moves by noise only
    assert np.allclose(a[ok].mean(axis=0), b[ok].mean(axis=0), rtol=0.05), (a[ok].mean(axis=0), b[ok].mean(axis=0))


@pytest.mark.gpu
@pytest.mark.parametrize("name,seed", [("spheres-1m", 3), ("many-lights", 5)])
def test_extension_bit_exact_at_baseline_size(name, seed):
    """BASELINE configs 3 and 5 (the two with smooth dielectrics) at 1080p, depth 8, every extension on: 2 frames of the
    production pipeline against the restatement on all host cores."""
    scene = m.Scene.procedural(name, seed)
    r, o = setup_pair(scene, 1920, 1080, seed=seed, max_path_length=8, renderer_kwargs=dict(pipeline=m.PIPELINE_WAVEFRONT))
    enable(r, o, scene, ALL, 0.02)
    tiles = m.noise_tiles(seed, 0, 2)
    r.draw(2, tiles); o.draw(2, tiles)
    gi, oi = r.save_current_image(), o.save_current_image()
    assert bits_equal(gi, oi), f"image differs on {np.mean(gi.view(np.uint32) != oi.view(np.uint32)):.3%} of the words"
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])


def _fuzz_cases():
    rng = np.random.default_rng(20260202)
    flag_sets = [m.EXT_REFRACTION, m.EXT_REFRACTION | m.EXT_DISPERSION, ALL, m.EXT_ALIAS_LIGHTS,
                 m.EXT_REFRACTION | m.EXT_ALIAS_LIGHTS, m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_CONTINUOUS_SPECTRUM]
    out = []
    for k in range(12):
        out.append(dict(scene=["spheres-64k", "many-lights", "cornell-water"][k % 3], w=int(rng.choice([3, 9, 33, 100, 129])),
                        h=int(rng.choice([2, 5, 17, 61])), L=int(rng.integers(1, 11)), frames=int(rng.integers(1, 6)),
                        first=int(rng.integers(0, 9)), stride=int(rng.integers(1, 4)), mode=int(rng.integers(0, 3)),
                        shard=(int(rng.integers(0, 2)), 2) if k % 4 == 3 else (0, 1), flags=flag_sets[k % len(flag_sets)],
                        spread=float(rng.choice([-0.05, 0.01, 0.1])), seed=100 + k))
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("c", _fuzz_cases(), ids=lambda c: f"{c['scene']}-{c['w']}x{c['h']}-L{c['L']}-f{c['frames']}-x{c['flags']}")
def test_extension_schedules_agree_on_odd_shapes(c):
    """Odd sizes, path length 1, skipped frames, row shards and the three accumulate modes with the extension on:
    wavefront == staged == restatement, bit for bit."""
    scene = water_as_glass() if c["scene"] == "cornell-water" else m.Scene.procedural(c["scene"], 1)
    span = (c["frames"] - 1) * c["stride"] + 1
    tiles = m.noise_tiles(c["seed"], c["first"], span)
    images = []
    for pipe in (m.PIPELINE_WAVEFRONT, m.PIPELINE_STAGED):
        r, o = setup_pair(scene, c["w"], c["h"], seed=c["seed"], max_path_length=c["L"],
                          renderer_kwargs=dict(pipeline=pipe, accumulate_mode=c["mode"]))
        enable(r, o, scene, c["flags"], c["spread"])
        r.set_row_shard(*c["shard"])
        r.draw(c["frames"], tiles, first_frame=c["first"], frame_stride=c["stride"])
        images.append(r.save_current_image())
        if pipe == m.PIPELINE_WAVEFRONT:
            o.set_accumulate_mode(c["mode"]); o.set_row_shard(*c["shard"])
            o.draw(c["frames"], tiles, first_frame=c["first"], frame_stride=c["stride"])
            assert_bits(images[0], o.save_current_image(), "wavefront vs restatement")
            st, oc = r.stats(), o.counts()
            assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])
    assert_bits(images[0], images[1], "wavefront vs staged")


# ---- white furnace: a physical check of the refraction lobe weights (CPU, on the restatement) -------------------------
def furnace_obj(path, rings=24, segments=48):
    """A closed box whose walls only emit (Kd 0, Ka 1) around the camera and a glass ball (ior 1.5, smooth normals)."""
    lines = ["mtllib furnace.mtl"]
    box = [(-1, 0, -1), (1, 0, -1), (1, 0, 3), (-1, 0, 3), (-1, 2, -1), (1, 2, -1), (1, 2, 3), (-1, 2, 3)]
    lines += ["v %g %g %g" % p for p in box]
    lines += ["usemtl wall"] + ["f %d %d %d %d" % q for q in [(1, 4, 3, 2), (5, 6, 7, 8), (1, 2, 6, 5), (2, 3, 7, 6), (3, 4, 8, 7), (4, 1, 5, 8)]]
    c, rad, first = np.array([0.0, 1.0, 0.0]), 0.5, len(box) + 1
    for i in range(rings + 1):
        th = math.pi * i / rings
        for j in range(segments):
            ph = 2 * math.pi * j / segments
            n = np.array([math.sin(th) * math.cos(ph), math.cos(th), math.sin(th) * math.sin(ph)])
            lines.append("v %.9g %.9g %.9g" % tuple(c + rad * n)); lines.append("vn %.9g %.9g %.9g" % tuple(n))
    lines.append("usemtl glass")
    idx = lambda i, j: i * segments + (j % segments)
    for i in range(rings):
        for j in range(segments):
            a, b, cc, d = idx(i, j), idx(i, j + 1), idx(i + 1, j + 1), idx(i + 1, j)
            for tri in ((a, b, cc), (a, cc, d)):                                   # counter-clockwise seen from outside
                lines.append("f " + " ".join("%d//%d" % (first + k, k + 1) for k in tri))
    path.write_text("\n".join(lines) + "\n")
    path.with_name("furnace.mtl").write_text("newmtl wall\nKd 0 0 0\nKa 1 1 1\nKs 1 0 0\nnewmtl glass\nKd 1 1 1\nKs 0 0 1.5\n")


def test_white_furnace_a_glass_ball_is_invisible_with_refraction_and_black_without(tmp_path):
    """Inside a box whose walls all emit L = 1 (and reflect nothing) every path that reaches a wall carries exactly the
    product of its lobe weights.  With MRT_EXT_REFRACTION both dielectric lobes weigh 1 and the eta^2 factors telescope
    to 1 once the path is back in air, so the ball must vanish: its pixels show L = 1.  The reference's branch leaves
    them black (the continued ray meets the same surface until its bounces are spent) — which is what the extension is for."""
    furnace_obj(tmp_path / "furnace.obj")
    scene = m.Scene.load_obj(str(tmp_path / "furnace.obj"))
    w, h, frames = 64, 48, 8
    tiles = m.noise_tiles(3, 0, frames)
    images = {}
    DISP = m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_CONTINUOUS_SPECTRUM
    for flags in (0, m.EXT_REFRACTION, DISP):
        o = orc.Oracle(max_path_length=8)
        o.init_raytracing(scene); o.drawable_size_will_change(w, h)
        for s in range(3):
            o.update_noise(s, m.noise_tile(3, 0xFFFFFFFF))
        o.set_extensions(flags, 0.03)
        o.raygen(0); o.intersect(0)
        first_hit = o.read_buffer(2)["triangleIndex"].reshape(h, w).copy()
        o.draw(frames, tiles)
        images[flags] = o.save_current_image()[..., :3]
    hit = first_hit >= 12                                        # 12 wall triangles come first (material order)
    grown, shrunk = hit.copy(), hit.copy()                       # frame 0's primary hits; later frames jitter by up to a pixel,
    for dy in (-1, 0, 1):                                        # so the silhouette band belongs to neither set
        for dx in (-1, 0, 1):
            moved = np.roll(np.roll(hit, dy, axis=0), dx, axis=1)
            grown |= moved; shrunk &= moved
    ball, walls = shrunk, ~grown
    assert 60 < ball.sum() < 1200 and walls.sum() > 2000
    assert np.allclose(images[0][walls], 1.0, atol=1e-6) and np.allclose(images[m.EXT_REFRACTION][walls], 1.0, atol=1e-6)
    seen = images[m.EXT_REFRACTION][ball]
    assert np.allclose(seen, 1.0, atol=1e-5), (seen.min(), seen.max())          # every pixel of the ball shows the wall's radiance
    assert images[0][ball].mean() < 0.2                          # the reference's straight-through branch: (nearly) black
    # dispersion keeps one channel with weight 3: each sample is (3,0,0), (0,3,0) or (0,0,3), the mean is still the wall's
    spectral = images[DISP][ball]
    assert np.allclose(spectral.sum(axis=1), 3.0, atol=1e-4) and np.allclose(spectral.mean(axis=0), 1.0, atol=0.1), spectral.mean(axis=0)
