"""Full-size GPU checks at BASELINE.json's sizes, through size-independent properties (the oracle would take
minutes here): wavefront == megakernel == the reference-shaped 34-dispatch sequence bit-for-bit, row shards compose,
traversal == brute-force definition on sampled rays, ray accounting."""
import numpy as np
import pytest

import mrt_b200 as m
from helpers import bits_equal, random_rays, setup_pair

pytestmark = pytest.mark.gpu


def gpu_renderer(scene, w, h, L=8, **kw):
    r = m.Renderer(max_path_length=L, **kw)
    r.init_raytracing(scene)
    r.drawable_size_will_change(w, h)
    init = m.noise_tile(1, 0xFFFFFFFF)
    for s in range(3):
        r.update_noise(s, init)
    return r


@pytest.mark.parametrize("name,seed", [("spheres-1m", 3), ("cornell-water", 2)])
def test_1080p_fused_equals_staged(name, seed):
    scene = m.Scene.procedural(name, seed)
    tiles = m.noise_tiles(seed, 0, 3)
    a = gpu_renderer(scene, 1920, 1080, pipeline=m.PIPELINE_WAVEFRONT)
    b = gpu_renderer(scene, 1920, 1080, pipeline=m.PIPELINE_STAGED)
    c = gpu_renderer(scene, 1920, 1080, pipeline=m.PIPELINE_MEGAKERNEL)
    a.draw(3, tiles); b.draw(3, tiles); c.draw(3, tiles)
    ia, ib, ic = a.save_current_image(), b.save_current_image(), c.save_current_image()
    assert bits_equal(ia, ib) and bits_equal(ic, ib)
    sa, sb, sc = a.stats(), b.stats(), c.stats()
    assert (sa["primary_rays"], sa["shadow_rays"]) == (sb["primary_rays"], sb["shadow_rays"]) == (sc["primary_rays"], sc["shadow_rays"])
    n = 1920 * 1080 * 3
    assert n <= sa["primary_rays"] <= 8 * n and 0 < sa["shadow_rays"] <= 7 * n
    finite = np.isfinite(ia[..., :3])
    assert finite.mean() > 0.9999 and (ia[..., 3] == 1).all()
    assert 0.005 < ia[..., :3][finite].mean() < 10


def test_1m_triangle_traversal_equals_oracle_on_samples():
    import oracle as orc
    scene = m.Scene.procedural("spheres-1m", 3)
    r = gpu_renderer(scene, 16, 16)
    o = orc.Oracle(); o.init_raytracing(scene)
    rays = random_rays(50000, 17)
    got = r.intersect_host(rays)
    want = o.intersect_host(rays, bvh=True)
    assert bits_equal(got, want)
    few = rays[:64]
    assert bits_equal(r.intersect_host(few), o.intersect_host(few, bvh=False))          # brute force over 1M triangles
    st = r.stats()
    assert st["bvh_depth"] < 62 and st["bvh_nodes"] > 200_000


def test_4k_row_shards_compose():
    scene = m.Scene.procedural("spheres-64k", 4)
    tiles = m.noise_tiles(4, 0, 2)
    full = gpu_renderer(scene, 3840, 2160); full.draw(2, tiles)
    ref = full.save_current_image()
    acc = np.zeros_like(ref)
    for g in range(2):
        r = gpu_renderer(scene, 3840, 2160); r.set_row_shard(g, 2); r.draw(2, tiles)
        acc[g::2] = r.save_current_image()[g::2]
        r.close()
    assert bits_equal(acc, ref)


_VARIANT_SCRIPT = r"""
import sys
sys.path[:0] = [%(pkg)r, %(tests)r]
import mrt_b200 as m
from helpers import bits_equal
from test_gpu_fullsize import gpu_renderer
scene = m.Scene.procedural("spheres-64k", 3)
tiles = m.noise_tiles(5, 0, 3)
a = gpu_renderer(scene, 1024, 576, pipeline=m.PIPELINE_WAVEFRONT)      # 589,824 slots: enough for two slot groups
b = gpu_renderer(scene, 1024, 576, pipeline=m.PIPELINE_MEGAKERNEL)     # binary tree, one thread per pixel
a.draw(3, tiles); b.draw(3, tiles)
sa, sb = a.stats(), b.stats()
assert bits_equal(a.save_current_image(), b.save_current_image()), "image differs"
assert (sa["primary_rays"], sa["shadow_rays"]) == (sb["primary_rays"], sb["shadow_rays"]), "ray counts differ"
print("variant ok", sa["kernel_launches"])
"""


@pytest.mark.parametrize("env", [
    {"MRT_WF_SPLIT": "2"}, {"MRT_WF_SPLIT": "3"},                       # slot groups on separate streams
    {"MRT_WF_REGION": "256"}, {"MRT_WF_REGION": "0"},                   # work-claim granularity
    {"MRT_BVH_WIDTH": "2"},                                             # binary tree in the persistent trace kernel
    {"MRT_TRACE_MINBLOCKS": "6", "MRT_TRACE_REFILL": "28", "MRT_TRACE_LEAFMIN": "4", "MRT_TRACE_TRIRATIO": "0"},
    {"MRT_SHADE_MINBLOCKS": "3"}, {"MRT_TRACE_COUNT": "1"},
    {"MRT_TRACE_DEFER": "1"}, {"MRT_TRACE_DEFER": "1", "MRT_TRACE_TRITHRESH": "28", "MRT_TRACE_LEAFMIN": "16"},   # postponed triangles
    {"MRT_WF_SLOTS": "2"}, {"MRT_WF_SLOTS": "3", "MRT_WF_SPLIT": "2"},  # more path slots than pixels (fewer, larger waves)
    {"MRT_BVH8_MAXLEAF": "1"}, {"MRT_BVH8_CPRIM": "0.3", "MRT_BVH8_FRONTLEAF": "3"},     # other shapes of the wide tree
    {"MRT_TRACE_TQ": "32"}, {"MRT_TRACE_TQ": "1"}, {"MRT_TRACE_TQ": "96", "MRT_TRACE_COUNT": "1"},   # warp-wide triangle queue
    {"MRT_BVH_BINS": "64"}, {"MRT_BVH_BINS": "8,32"},                   # SAH bins of the binary builder: fixed / other candidates for the sampled selection
    {"MRT_FUSED_BVH": "8"}, {"MRT_FUSED_BVH": "8", "MRT_FUSED_MINBLOCKS": "2"},   # megakernel on the wide traversal core (default: binary tree)
    {"MRT_WF_GRAPH": "0"}, {"MRT_WF_GRAPH": "0", "MRT_WF_SORT": "2"},   # polling host loop instead of the WHILE graph
    {"MRT_WF_SORT": "0"}, {"MRT_WF_SORT": "1"}, {"MRT_WF_SORT": "2", "MRT_WF_SORT_MODE": "1"}, {"MRT_WF_SORT": "3"},   # ray ordering grid
    {"MRT_WF_SORT": "2", "MRT_WF_SPLIT": "2"}, {"MRT_WF_SORT": "3", "MRT_TRACE_TQ": "32"},
], ids=lambda e: ",".join(f"{k[4:].lower()}={v}" for k, v in e.items()))
def test_tuning_switches_do_not_change_the_bits(env):
    """Every scheduling knob of csrc/wavefront.cu / api.cu / bvh.cpp (read once per process from the environment) must
    leave the image and the ray counts bit-identical: the wavefront render under the knob == the megakernel render.
    Runs against the EXPERIMENTS build (lib/libmrt_b200_exp.so, `make exp`): the production library compiles the
    rejected kernel variants and most of their knobs out."""
    import os, pathlib, subprocess, sys
    root = pathlib.Path(__file__).resolve().parent.parent
    exp = root / "metal-renderer_b200" / "lib" / "libmrt_b200_exp.so"
    if not exp.exists():
        subprocess.check_call(["make", "-s", "-C", str(root / "metal-renderer_b200"), "lib/libmrt_b200_exp.so"])
    script = _VARIANT_SCRIPT % dict(pkg=str(root / "metal-renderer_b200"), tests=str(root / "tests"))
    out = subprocess.run([sys.executable, "-c", script], env=dict(os.environ, MRT_B200_LIB=str(exp), **env), capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "variant ok" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_a_short_schedule_is_reported_not_returned():
    """Failure detection: if the wavefront schedule ends with rays still queued (here forced with MRT_WF_MAX_WAVES), the
    next call that hands results to the caller fails with MRT_E_STATE instead of returning a partial image."""
    import os, pathlib, subprocess, sys
    root = pathlib.Path(__file__).resolve().parent.parent
    script = r"""
import sys
sys.path[:0] = [%r, %r]
import mrt_b200 as m
from test_gpu_fullsize import gpu_renderer
r = gpu_renderer(m.Scene.procedural("cornell", 0), 128, 96, pipeline=m.PIPELINE_WAVEFRONT)
try:
    r.draw(4, m.noise_tiles(1, 0, 4))
    r.save_current_image()
    print("no error")
except m.MrtError as e:
    print("error", e.code)
""" % (str(root / "metal-renderer_b200"), str(root / "tests"))
    out = subprocess.run([sys.executable, "-c", script], env=dict(os.environ, MRT_WF_MAX_WAVES="3"), capture_output=True, text=True, timeout=300)
    assert "error -5" in out.stdout, out.stdout[-1000:] + out.stderr[-1000:]
