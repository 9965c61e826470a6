"""Seeded sweep over odd image sizes, path lengths, batch sizes, row shards and accumulate modes: the three GPU schedules
(wavefront with dynamic work items, megakernel, the reference-shaped staged sequence) must agree bit for bit on every
one, and a sample of them is also held against the CPU oracle.  Sizes that are not multiples of the 8x4 tile, a single
row of tiles, one-frame batches and path length 1 are the cases the work-item / padding-pixel logic has to survive."""
import numpy as np
import pytest

import mrt_b200 as m
from helpers import bits_equal, setup_pair

pytestmark = pytest.mark.gpu

SCENES = ["cornell", "cornell-water", "white-box", "many-lights", "spheres-64k"]


def _cases():
    rng = np.random.default_rng(20260101)
    out = []
    for k in range(36):
        w = int(rng.choice([2, 3, 7, 8, 9, 31, 33, 64, 100, 129, 255, 320]))
        h = int(rng.choice([2, 3, 4, 5, 9, 17, 32, 61, 128, 200]))
        out.append(dict(scene=SCENES[k % len(SCENES)], w=w, h=h, L=int(rng.integers(1, 11)), frames=int(rng.integers(1, 7)),
                        first=int(rng.integers(0, 9)), stride=int(rng.integers(1, 4)), mode=int(rng.integers(0, 3)),
                        shard=(int(rng.integers(0, 3)), 3) if k % 4 == 3 else (0, 1), oracle=(k % 6 == 0), seed=k))
    return out


@pytest.mark.parametrize("c", _cases(), ids=lambda c: f"{c['scene']}-{c['w']}x{c['h']}-L{c['L']}-f{c['frames']}")
def test_schedules_agree(c):
    scene = m.Scene.procedural(c["scene"], 1)
    span = (c["frames"] - 1) * c["stride"] + 1
    tiles = m.noise_tiles(c["seed"], c["first"], span)
    images, counts = [], []
    for pipe in (m.PIPELINE_WAVEFRONT, m.PIPELINE_MEGAKERNEL, m.PIPELINE_STAGED):
        r, o = setup_pair(scene, c["w"], c["h"], seed=c["seed"], max_path_length=c["L"],
                          renderer_kwargs=dict(pipeline=pipe, accumulate_mode=c["mode"]))
        r.set_row_shard(*c["shard"])
        r.draw(c["frames"], tiles, first_frame=c["first"], frame_stride=c["stride"])
        images.append(r.save_current_image())
        st = r.stats()
        counts.append((st["primary_rays"], st["shadow_rays"]))
        if c["oracle"] and pipe == m.PIPELINE_WAVEFRONT:
            o.set_accumulate_mode(c["mode"]); o.set_row_shard(*c["shard"])
            o.draw(c["frames"], tiles, first_frame=c["first"], frame_stride=c["stride"])
            assert bits_equal(images[0], o.save_current_image()), "wavefront vs oracle"
            oc = o.counts()
            assert counts[0] == (oc["primary_rays"], oc["shadow_rays"])
    assert bits_equal(images[0], images[2]), "wavefront vs staged"
    assert bits_equal(images[1], images[2]), "megakernel vs staged"
    assert counts[0] == counts[1] == counts[2]
