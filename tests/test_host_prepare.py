"""Host half of the scene setup (mrt_scene_prepare: validation + both SAH trees, no GPU needed).  Covers the input
checks mrt_set_scene relies on — among them the one ADVICE r1 asked for: the kernels pick a light by bisection over the
cdf, which equals the reference's linear scan (KernelHelpers.h:49-54) only if the cdf never decreases."""
import ctypes as C

import numpy as np
import pytest

import mrt_b200 as m
from mrt_b200 import _lib as L


def prepare(scene, lights=None, light_count=None, references=None, indices=None):
    lights = scene.lights if lights is None else lights
    references = scene.references if references is None else references
    indices = scene.indices if indices is None else indices
    out = C.c_void_p()
    rc = L.lib().mrt_scene_prepare(L.ptr(scene.vertices), len(scene.vertices), L.ptr(indices), len(references), L.ptr(references),
                                   L.ptr(scene.materials), len(scene.materials), L.ptr(lights),
                                   scene.light_count if light_count is None else light_count, 4, 2, C.byref(out))
    if rc == 0:
        L.lib().mrt_prepared_scene_destroy(out)
    return rc, L.lib().mrt_last_error().decode()


def test_prepare_accepts_the_procedural_scenes():
    for name in ("cornell", "cornell-water", "many-lights"):
        rc, msg = prepare(m.Scene.procedural(name, 1))
        assert rc == 0, msg


def test_prepare_rejects_a_decreasing_light_cdf():
    s = m.Scene.procedural("many-lights", 5)
    lights = s.lights.copy()
    lights["cdf"][7], lights["cdf"][8] = lights["cdf"][8], lights["cdf"][7]
    assert lights["cdf"][7] > lights["cdf"][8]
    rc, msg = prepare(s, lights=lights)
    assert rc == -1 and "cdf decreases at 7" in msg


def test_prepare_rejects_out_of_range_indices():
    s = m.Scene.procedural("cornell", 0)
    idx = s.indices.copy(); idx[5] = len(s.vertices)
    rc, msg = prepare(s, indices=idx)
    assert rc == -1 and "out of range" in msg
    refs = s.references.copy(); refs["materialIndex"][3] = len(s.materials)
    rc, msg = prepare(s, references=refs)
    assert rc == -1 and "materialIndex" in msg
    refs = s.references.copy(); refs["lightTriangleIndex"][0] = s.light_count
    rc, msg = prepare(s, references=refs)
    assert rc == -1 and "lightTriangleIndex" in msg
    lights = s.lights.copy(); lights["index"][0] = s.triangle_count
    rc, msg = prepare(s, lights=lights)
    assert rc == -1 and "light 0" in msg


def test_wide_tree_does_not_depend_on_the_thread_count():
    """The host build is parallel in its front, collapse and emission; the emitted nodes and triangle records (FNV hashes
    printed under MRT_BVH_VERBOSE) must be the same for 1 thread, 3 threads and all cores."""
    import os, pathlib, re, subprocess, sys
    root = pathlib.Path(__file__).resolve().parent.parent
    script = r"""
import sys, ctypes as C
sys.path[:0] = [%r]
import mrt_b200 as m
from mrt_b200 import _lib as L
s = m.Scene.procedural("spheres-64k", 3)
out = C.c_void_p()
rc = L.lib().mrt_scene_prepare(L.ptr(s.vertices), len(s.vertices), L.ptr(s.indices), len(s.references), L.ptr(s.references),
                               L.ptr(s.materials), len(s.materials), L.ptr(s.lights), s.light_count, 4, int(sys.argv[1]), C.byref(out))
assert rc == 0
""" % str(root / "metal-renderer_b200")
    seen = []
    for threads in (1, 3, 0):
        run = subprocess.run([sys.executable, "-c", script, str(threads)], env=dict(os.environ, MRT_BVH_VERBOSE="1"),
                             capture_output=True, text=True, timeout=300)
        assert run.returncode == 0, run.stderr[-1500:]
        hashes = re.findall(r"\((\d+) bins\): (\d+) nodes ([0-9a-f]{16}), (\d+) triangle records ([0-9a-f]{16}), depth (\d+)", run.stderr)
        assert len(hashes) == 2, run.stderr[-1500:]                    # the two SAH-bin candidates
        seen.append(hashes)
    assert seen[0] == seen[1] == seen[2], seen
