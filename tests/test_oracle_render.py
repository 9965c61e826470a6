"""CPU-side checks of the oracle's intersector and frame loop (no GPU): BVH == brute force bit-for-bit,
stage-by-stage sequence == fused frame loop, energy sanity on the furnace-like white box."""
import numpy as np

import mrt_b200 as m
import oracle as orc
from helpers import bits_equal, random_rays


def make_oracle(scene, w, h, seed=1, **kw):
    o = orc.Oracle(**kw)
    o.init_raytracing(scene)
    o.drawable_size_will_change(w, h)
    init = m.noise_tile(seed, 0xFFFFFFFF)
    for s in range(3):
        o.update_noise(s, init)
    return o


def test_oracle_bvh_equals_brute_force():
    for name in ("cornell", "cornell-water"):
        scene = m.Scene.procedural(name, 2)
        o = orc.Oracle(threads=0)
        o.init_raytracing(scene)
        rays = random_rays(4000 if name == "cornell" else 1500, 7)
        a = o.intersect_host(rays, bvh=True)
        b = o.intersect_host(rays, bvh=False)
        assert bits_equal(a, b)
        assert (a["distance"] >= 0).mean() > 0.5


def test_dead_rays_are_skipped_and_miss_is_negative():
    scene = m.Scene.procedural("cornell", 0)
    o = orc.Oracle(threads=1); o.init_raytracing(scene)
    rays = random_rays(64, 3)
    rays["maxDistance"][::2] = -1.0
    out = o.intersect_host(rays)
    assert (out["distance"][::2] == -1.0).all() and (out["triangleIndex"][::2] == 0xFFFFFFFF).all()
    up = random_rays(4, 1); up["origin"] = [0, 1, 5]; up["direction"] = [0, 0, 1]      # away from the box
    assert (o.intersect_host(up)["distance"] < 0).all()


def test_staged_sequence_equals_frame_loop():
    scene = m.Scene.procedural("cornell", 0)
    w, h, L = 48, 32, 4
    tiles = m.noise_tiles(5, 0, 3)
    a = make_oracle(scene, w, h, max_path_length=L, threads=1)
    b = make_oracle(scene, w, h, max_path_length=L, threads=3)
    for f in range(3):                                   # Renderer.mm:472-498 then 500-585, dispatch by dispatch
        a.update_noise(f % 3, tiles[f])
        a.raygen(f)
        for i in range(L):
            a.intersect(0); a.shade(f, i); a.intersect(1); a.shadow_resolve()
        a.accumulate(f)
    b.draw(3, tiles)
    assert bits_equal(a.save_current_image(), b.save_current_image())
    assert bits_equal(a.read_buffer(0), b.read_buffer(0))
    assert a.counts()["primary_rays"] == b.counts()["primary_rays"] > 0


def test_white_box_energy_sanity():
    # all-white closed-ish box lit by a pi-radiance light: image must be finite, positive, and brighter with depth
    scene = m.Scene.procedural("white-box", 0)
    means = []
    for L in (2, 6):
        o = make_oracle(scene, 64, 48, max_path_length=L)
        o.draw(8, m.noise_tiles(9, 0, 8))
        img = o.save_current_image()
        assert np.isfinite(img).all() and (img[..., 3] == 1).all()
        means.append(img[..., :3].mean())
    assert 0.01 < means[0] < means[1] < 5.0


def test_row_shards_compose_to_full_image():
    scene = m.Scene.procedural("cornell", 0)
    w, h = 40, 30
    tiles = m.noise_tiles(4, 0, 2)
    full = make_oracle(scene, w, h, max_path_length=3); full.draw(2, tiles)
    ref = full.save_current_image()
    acc = np.zeros_like(ref)
    for g in range(3):
        o = make_oracle(scene, w, h, max_path_length=3); o.set_row_shard(g, 3); o.draw(2, tiles)
        img = o.save_current_image()
        acc[g::3] = img[g::3]
        assert (img[(g + 1) % 3::3] == 0).all()
    assert bits_equal(acc, ref)
