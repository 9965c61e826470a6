"""GPU parity tests: the CUDA path (through the C-ABI of include/mrt.h) against the CPU oracle on the same seeded
inputs.  Bar: BIT-EXACT for hit IDs, distances, barycentrics, every stage buffer and the accumulated image
(NaNs compare equal to NaNs).  Sizes here are ones the oracle finishes in seconds."""
import numpy as np
import pytest

import mrt_b200 as m
from helpers import bits_equal, mismatch_report, random_rays, setup_pair

pytestmark = pytest.mark.gpu


def assert_bits(a, b, what):
    assert bits_equal(a, b), f"{what}: {mismatch_report(a, b) if a.dtype.fields else 'arrays differ'}"


def test_raygen_buffer_bits():
    scene = m.Scene.procedural("cornell", 0)
    for (w, h) in ((256, 256), (101, 37), (2, 2)):
        r, o = setup_pair(scene, w, h)
        r.raygen(0); o.raygen(0)
        assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"rays {w}x{h}")


@pytest.mark.parametrize("name,n", [("cornell", 20000), ("cornell-water", 20000), ("spheres-64k", 20000), ("many-lights", 8000)])
def test_intersector_matches_oracle(name, n):
    scene = m.Scene.procedural(name, 3)
    r, o = setup_pair(scene, 8, 8)
    rays = random_rays(n, 11)
    rays["maxDistance"][::17] = -1.0                       # skipped rays
    rays["maxDistance"][1::19] = 0.7                       # finite range
    rays["minDistance"][2::23] = 0.3
    # axis-parallel directions (exact zero components): 1/d is infinite — the slab test must stay conservative
    axes = np.array([[0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 0.6, -0.8], [-0.6, 0, 0.8], [0.8, -0.6, 0], [-0.0, 1, 0]], np.float32)
    rays["direction"][3::29] = axes[np.arange(len(rays["direction"][3::29])) % len(axes)]
    rays["origin"][3::58] = np.round(rays["origin"][3::58] * 4) / 4     # some origins exactly on wall / box planes
    got = r.intersect_host(rays)
    want = o.intersect_host(rays, bvh=(name != "cornell"))  # brute force defines the answer on the small scene
    assert_bits(got, want, name)
    # strided records (80-byte Ray layout, MPS rayStride) give the same answer
    wide = random_rays(n, 11, dtype=m.RAY)
    for f in ("origin", "direction", "maxDistance", "minDistance"):
        wide[f] = rays[f]
    assert_bits(r.intersect_host(wide), want, name + " stride 80")


def test_every_stage_buffer_matches_oracle():
    scene = m.Scene.procedural("cornell-water", 0)         # diffuse + plastic + mirror
    w, h, L = 96, 64, 5
    r, o = setup_pair(scene, w, h, max_path_length=L)
    tiles = m.noise_tiles(21, 0, 3)
    for f in range(3):
        r.update_noise(f % 3, tiles[f]); o.update_noise(f % 3, tiles[f])
        r.raygen(f); o.raygen(f)
        assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"f{f} raygen")
        for i in range(L):
            r.intersect(m.BUF_RAYS); o.intersect(0)
            assert_bits(r.read_buffer(m.BUF_INTERSECTIONS), o.read_buffer(2), f"f{f} i{i} primary hits")
            r.shade(f, i); o.shade(f, i)
            assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"f{f} i{i} shade rays")
            assert_bits(r.read_buffer(m.BUF_LIGHT_RAYS), o.read_buffer(1), f"f{f} i{i} light rays")
            r.intersect(m.BUF_LIGHT_RAYS); o.intersect(1)
            assert_bits(r.read_buffer(m.BUF_INTERSECTIONS), o.read_buffer(2), f"f{f} i{i} shadow hits")
            r.shadow_resolve(); o.shadow_resolve()
            assert_bits(r.read_buffer(m.BUF_RAYS), o.read_buffer(0), f"f{f} i{i} resolve")
        r.accumulate(f); o.accumulate(f)
        assert_bits(r.read_buffer(m.BUF_IMAGE), o.read_buffer(3), f"f{f} image")


def render_both(scene, w, h, frames, L, seed, pipeline=m.PIPELINE_AUTO, **kw):
    r, o = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=pipeline))
    tiles = m.noise_tiles(seed, 0, frames)
    r.draw(frames, tiles); o.draw(frames, tiles)
    return r, o


PIPES = [pytest.param(m.PIPELINE_WAVEFRONT, id="wavefront"), pytest.param(m.PIPELINE_MEGAKERNEL, id="megakernel")]


@pytest.mark.parametrize("pipeline", PIPES)
def test_config1_cornell_256_16spp_depth4_bit_exact(pipeline):
    """BASELINE.json configs[0]: procedural Cornell 256x256, 16 spp, max depth 4."""
    scene = m.Scene.procedural("cornell", 0)
    r, o = render_both(scene, 256, 256, 16, 4, seed=1, pipeline=pipeline)
    gi, oi = r.save_current_image(), o.save_current_image()
    assert_bits(gi, oi, "image")
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])
    if pipeline == m.PIPELINE_MEGAKERNEL:
        assert st["kernel_launches"] == 1
    else:       # first shade + resolve + (trace, shade) per wave; the WHILE graph stops within two launches of the last
        # wave that had rays, far below the static bound of 16 * 4 + 4 + 2 waves
        assert 2 + 2 * st["waves_traced"] <= st["kernel_launches"] <= 2 + 2 * (st["waves_traced"] + 3) and 16 <= st["waves_traced"] <= 70
    # primary-hit triangle IDs: bit-equal on 100% of pixels
    r2, o2 = setup_pair(scene, 256, 256)
    r2.raygen(0); o2.raygen(0); r2.intersect(); o2.intersect()
    assert np.array_equal(r2.read_buffer(m.BUF_INTERSECTIONS)["triangleIndex"], o2.read_buffer(2)["triangleIndex"])


@pytest.mark.parametrize("name,w,h,frames,L", [("cornell-water", 160, 90, 6, 8), ("spheres-64k", 128, 72, 4, 8),
                                                ("many-lights", 96, 54, 4, 8), ("white-box", 64, 64, 5, 3)])
@pytest.mark.parametrize("pipeline", PIPES)
def test_production_render_bit_exact(name, w, h, frames, L, pipeline):
    scene = m.Scene.procedural(name, 5)
    r, o = render_both(scene, w, h, frames, L, seed=7, pipeline=pipeline)
    assert_bits(r.save_current_image(), o.save_current_image(), name)
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])


def test_staged_pipeline_and_continuation_bit_exact():
    scene = m.Scene.procedural("cornell-water", 0)
    w, h, L = 120, 68, 6
    tiles = m.noise_tiles(3, 0, 7)
    a, o = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=m.PIPELINE_WAVEFRONT))
    c, _ = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=m.PIPELINE_MEGAKERNEL))
    b, _ = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=m.PIPELINE_STAGED))
    a.draw(3, tiles[:3]); a.draw(4, tiles[3:])                                           # noise slots carry over between calls
    c.draw(3, tiles[:3]); c.draw(4, tiles[3:])
    b.draw(7, tiles)
    o.draw(7, tiles)
    want = o.save_current_image()
    assert_bits(a.save_current_image(), want, "wavefront 3+4")
    assert_bits(c.save_current_image(), want, "megakernel 3+4")
    assert_bits(b.save_current_image(), want, "staged")
    assert b.stats()["kernel_launches"] == 7 * (2 + 4 * L)
    sa, sb, sc = a.stats(), b.stats(), c.stats()
    assert (sa["primary_rays"], sa["shadow_rays"]) == (sb["primary_rays"], sb["shadow_rays"]) == (sc["primary_rays"], sc["shadow_rays"])


@pytest.mark.parametrize("pipeline", PIPES)
def test_noise_zero_produces_identical_nans(pipeline):
    # noise.w == 0 makes bsdf/pdf = 0/0 (Shaders.metal:209): the NaNs must appear in the same pixels
    scene = m.Scene.procedural("cornell", 0)
    tiles = m.noise_tiles(2, 0, 3).copy()
    tiles[:, 3::4][:, ::5] = 0.0
    tiles[:, 2::4][:, ::7] = 0.0
    r, o = setup_pair(scene, 64, 64, max_path_length=4, renderer_kwargs=dict(pipeline=pipeline))
    r.draw(3, tiles); o.draw(3, tiles)
    gi, oi = r.save_current_image(), o.save_current_image()
    assert_bits(gi, oi, "image with NaNs")


@pytest.mark.parametrize("pipeline", PIPES)
def test_accumulate_modes_and_row_shards(pipeline):
    scene = m.Scene.procedural("cornell", 0)
    w, h, L = 64, 48, 4
    tiles = m.noise_tiles(8, 0, 4)
    for mode in (m.ACCUM_SUM, m.ACCUM_REPLACE):
        r, o = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(accumulate_mode=mode, pipeline=pipeline))
        o.set_accumulate_mode(mode)
        r.draw(4, tiles); o.draw(4, tiles)
        assert_bits(r.save_current_image(), o.save_current_image(), f"mode {mode}")
    # row sharding: 3 shards compose to the unsharded image bit-for-bit; strided frames match the oracle
    full, o = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=pipeline))
    full.draw(4, tiles)
    ref = full.save_current_image()
    acc = np.zeros_like(ref)
    for g in range(3):
        r, _ = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(pipeline=pipeline))
        r.set_row_shard(g, 3)
        r.draw(4, tiles)
        acc[g::3] = r.save_current_image()[g::3]
    assert_bits(acc, ref, "row shards")
    r, o = setup_pair(scene, w, h, max_path_length=L, renderer_kwargs=dict(accumulate_mode=m.ACCUM_SUM, pipeline=pipeline))
    o.set_accumulate_mode(m.ACCUM_SUM)
    r.draw(2, tiles, first_frame=1, frame_stride=2); o.draw(2, tiles, first_frame=1, frame_stride=2)   # frames 1, 3
    assert_bits(r.save_current_image(), o.save_current_image(), "strided frames")


def test_error_paths():
    r = m.Renderer()
    with pytest.raises(m.MrtError) as e:
        r.raygen(0)
    assert e.value.code == -5                                  # MRT_E_STATE: no scene
    scene = m.Scene.procedural("cornell", 0)
    r.init_raytracing(scene)
    with pytest.raises(m.MrtError):
        r.drawable_size_will_change(1, 10)                     # raygen divides by size-1
    bad = m.Scene.procedural("cornell", 0)
    bad.indices[5] = 10_000
    with pytest.raises(m.MrtError) as e:
        r.init_raytracing(bad)
    assert e.value.code == -1
    with pytest.raises(m.MrtError):
        m.Renderer(max_path_length=0)


def test_empty_and_single_triangle_scene():
    import oracle as orc
    scene = m.Scene.procedural("cornell", 0)
    one = m.Scene(scene.vertices, scene.indices[:3], scene.references[:1], scene.materials, scene.lights[-1:], 0)
    r, o = setup_pair(one, 32, 32)
    rays = random_rays(2000, 5)
    assert_bits(r.intersect_host(rays), o.intersect_host(rays, bvh=False), "one triangle")
    r.draw(2, m.noise_tiles(1, 0, 2)); o.draw(2, m.noise_tiles(1, 0, 2))
    assert_bits(r.save_current_image(), o.save_current_image(), "one triangle image")


def test_rays_starting_on_shared_edges_and_vertices():
    """Hits at t = +0 / -0 on two coplanar triangles at once (origin on their shared edge, or on a vertex shared by many):
    the tie goes to the lowest triangle index, and the nearest-hit bound becoming a signed zero must not cull the other
    candidate (the device node test works on the sign of tf - tn, csrc/bvh8.h)."""
    for name in ("cornell", "spheres-64k"):
        scene = m.Scene.procedural(name, 3)
        r, o = setup_pair(scene, 8, 8)
        idx = scene.indices.reshape(-1, 3)[:400]
        v = scene.vertices["v"]
        origins = np.concatenate([v[idx[:, 0]], 0.5 * (v[idx[:, 0]] + v[idx[:, 1]]), 0.5 * (v[idx[:, 1]] + v[idx[:, 2]]),
                                  (v[idx[:, 0]] + v[idx[:, 1]] + v[idx[:, 2]]) / np.float32(3)]).astype(np.float32)
        dirs = np.array([[0, -1, 0], [0, 1, 0], [1, 0, 0], [-1, 0, 0], [0, 0, 1], [0, 0, -1], [0.6, -0.8, 0], [0, 0.6, 0.8],
                         [-0.57735, 0.57735, -0.57735]], np.float32)
        rays = random_rays(len(origins) * len(dirs), 1)
        rays["origin"] = np.repeat(origins, len(dirs), axis=0)
        rays["direction"] = np.tile(dirs, (len(origins), 1))
        rays["maxDistance"][::7] = -0.0                      # a signed-zero range is a live ray (not < 0)
        rays["maxDistance"][3::11] = 0.0
        got = r.intersect_host(rays)
        want = o.intersect_host(rays, bvh=(name != "cornell"))
        assert_bits(got, want, name)
        assert (want["distance"] == 0).sum() > 50            # the case is actually exercised


@pytest.mark.gpu
def test_visit_counters_leave_the_image_alone():
    """mrt_set_profile(MRT_PROFILE_VISITS): the counting variant of the trace kernel (the BVH term of bench.py's roofline)
    renders the same bits and reports node / triangle tests; without the flag the counters stay at zero."""
    scene = m.Scene.procedural("spheres-64k", 5)
    tiles = m.noise_tiles(6, 2, 4)
    plain, o = setup_pair(scene, 120, 72, max_path_length=6)
    plain.draw(4, tiles); o.draw(4, tiles)
    assert bits_equal(plain.save_current_image(), o.save_current_image())
    s0 = plain.stats()
    assert s0["nodes_visited"] == 0 and s0["triangles_tested"] == 0
    counted, _ = setup_pair(scene, 120, 72, max_path_length=6)
    counted.set_profile(m.PROFILE_VISITS | m.PROFILE_TRACE_TIME)
    counted.draw(4, tiles)
    assert bits_equal(counted.save_current_image(), plain.save_current_image())
    s1 = counted.stats()
    rays = s1["primary_rays"] + s1["shadow_rays"]
    assert rays == s0["primary_rays"] + s0["shadow_rays"]
    assert s1["nodes_visited"] >= rays                       # every traced ray tests at least the root
    assert 0 < s1["triangles_tested"] < 64 * rays
    assert s1["profile_trace_launches"] > 0


def test_host_loop_schedule_equals_the_graph_schedule_and_stops_early():
    """mrt_set_wave_schedule: the WHILE graph and the polling host loop run the same kernels — same image, same ray
    counts — and neither launches the static bound of frames * L + L + 2 waves when the paths die early."""
    scene = m.Scene.procedural("cornell-water", 1)
    w, h, frames, L = 320, 180, 24, 8
    tiles = m.noise_tiles(11, 0, frames)
    out = []
    for schedule in (m.SCHEDULE_GRAPH, m.SCHEDULE_HOST_LOOP):
        r, _ = setup_pair(scene, w, h, seed=11, max_path_length=L)
        r.set_wave_schedule(schedule)
        r.draw(frames, tiles)
        st = r.stats()
        out.append((r.save_current_image(), st))
        waves, launches = st["waves_traced"], st["kernel_launches"]
        assert 2 + 2 * waves <= launches, (schedule, waves, launches)
        slack = 3 if schedule == m.SCHEDULE_GRAPH else 26            # graph: ends inside the wave pair; loop: polls every 8 waves, 16 ahead
        assert launches <= 2 + 2 * (waves + slack), (schedule, waves, launches)
        assert waves < frames * L + L + 2
    assert_bits(out[0][0], out[1][0], "graph vs host loop")
    assert (out[0][1]["primary_rays"], out[0][1]["shadow_rays"]) == (out[1][1]["primary_rays"], out[1][1]["shadow_rays"])
