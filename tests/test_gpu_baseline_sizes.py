"""Parity at BASELINE.json's own sizes: every config rendered by the production (wavefront) pipeline AND by the oracle
on all host cores at the config's resolution and path length — accumulated image bit for bit, live-ray counts, and the
primary-hit triangle index of 100 % of the pixels (raygen + intersect on both sides).  Frames are kept to what the
oracle renders in about a minute (it does 1080p over 1 M triangles in ~1.4 s per frame on 16 cores); the bench line
repeats the same gate on the benchmarked workload (bench.py: `parity`)."""
import pathlib

import numpy as np
import pytest

import mrt_b200 as m
from helpers import bits_equal, setup_pair

pytestmark = pytest.mark.gpu

CONFIGS = [   # BASELINE.json configs[1..4]; SURVEY.md 8(d) "Concrete inputs"
    pytest.param("media:CornellBox-Water-plastic", 2, 1920, 1080, 8, 2, id="config2-media-water-plastic-1080p"),   # the reference's own file
    pytest.param("cornell-water", 2, 1920, 1080, 8, 2, id="config2-cornell-water-1080p"),
    pytest.param("spheres-1m", 3, 1920, 1080, 8, 2, id="config3-spheres-1m-1080p"),
    pytest.param("spheres-10m", 4, 3840, 2160, 8, 1, id="config4-spheres-10m-4k"),
    pytest.param("many-lights", 5, 1920, 1080, 8, 2, id="config5-many-lights-1080p"),
]


@pytest.mark.parametrize("name,seed,w,h,depth,frames", CONFIGS)
def test_baseline_config_bit_exact_at_its_own_size(name, seed, w, h, depth, frames):
    if name.startswith("media:"):       # renderer/Media/<name>.obj through the OBJ/MTL loader, committed as ABI arrays
        scene = m.Scene.from_npz(pathlib.Path(__file__).resolve().parent / "golden" / "media_water.npz", name[6:])
    else:
        scene = m.Scene.procedural(name, seed)
    r, o = setup_pair(scene, w, h, seed=seed, max_path_length=depth, renderer_kwargs=dict(pipeline=m.PIPELINE_WAVEFRONT))
    # primary hits of frame 0 through the staged entry points (the drop-in seam): 100 % of the pixels
    r.raygen(0); o.raygen(0)
    assert bits_equal(r.read_buffer(m.BUF_RAYS), o.read_buffer(0)), "primary rays"
    r.intersect(m.BUF_RAYS); o.intersect(0)
    gh, oh = r.read_buffer(m.BUF_INTERSECTIONS), o.read_buffer(2)
    agree = float(np.mean(gh["triangleIndex"] == oh["triangleIndex"]))
    assert agree == 1.0, f"primary-hit triangle ids agree on {agree:.6%} of {w * h} pixels"
    assert bits_equal(gh, oh), "distance / barycentrics of the primary hits"
    assert (gh["distance"] >= 0).mean() > 0.5                       # the camera looks into the box
    # production render
    tiles = m.noise_tiles(seed, 0, frames)
    r.reset_stats(); o.reset_counts()
    r.draw(frames, tiles); o.draw(frames, tiles)
    gi, oi = r.save_current_image(), o.save_current_image()
    assert bits_equal(gi, oi), f"image differs on {np.mean(gi.view(np.uint32) != oi.view(np.uint32)):.3%} of the words"
    st, oc = r.stats(), o.counts()
    assert (st["primary_rays"], st["shadow_rays"]) == (oc["primary_rays"], oc["shadow_rays"])
    assert st["primary_rays"] >= w * h * frames
    fin = np.isfinite(gi[..., :3])
    assert fin.mean() > 0.9999 and 0.003 < gi[..., :3][fin].mean() < 20
