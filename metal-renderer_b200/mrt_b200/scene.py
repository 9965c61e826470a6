"""Host-side scenes in the reference's buffer ABI (caller side of the path: what -initRaytracing assembles,
reference renderer/Renderer.mm:257-448).  Backed by the C++ generators / OBJ loader of libmrt_b200.so."""
import ctypes as C

import numpy as np

from . import _lib as L

PROCEDURAL_SCENES = ("cornell", "white-box", "cornell-water", "spheres-64k", "spheres-1m", "spheres-10m", "many-lights")


class Scene:
    """Arrays: vertices (VERTEX), indices (uint32, 3/triangle), references (TRIANGLE_REFERENCE),
    materials (MATERIAL), lights (LIGHT_TRIANGLE, light_count + 1 sentinel)."""

    def __init__(self, vertices, indices, references, materials, lights, light_count, name="custom"):
        self.vertices = np.ascontiguousarray(vertices, dtype=L.VERTEX)
        self.indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        self.references = np.ascontiguousarray(references, dtype=L.TRIANGLE_REFERENCE)
        self.materials = np.ascontiguousarray(materials, dtype=L.MATERIAL)
        self.lights = np.ascontiguousarray(lights, dtype=L.LIGHT_TRIANGLE)
        self.light_count = int(light_count)
        self.name = name
        assert len(self.lights) == self.light_count + 1, "light array must carry the sentinel entry"
        assert len(self.indices) == 3 * len(self.references)

    @property
    def triangle_count(self):
        return len(self.references)

    @classmethod
    def _from_handle(cls, handle, name):
        view = L.SceneView()
        L.check(L.lib().mrt_scene_get(handle, C.byref(view)))

        def arr(p, count, dtype):
            if count == 0:
                return np.zeros(0, dtype=dtype)
            buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(p)
            return np.frombuffer(buf, dtype=dtype).copy()

        scene = cls(arr(view.vertices, view.vertex_count, L.VERTEX),
                    arr(view.indices, 3 * view.triangle_count, np.uint32),
                    arr(view.references, view.triangle_count, L.TRIANGLE_REFERENCE),
                    arr(view.materials, view.material_count, L.MATERIAL),
                    arr(view.lights, view.light_count + 1, L.LIGHT_TRIANGLE), view.light_count, name)
        L.lib().mrt_scene_destroy(handle)
        return scene

    @classmethod
    def procedural(cls, name, seed=0):
        handle = C.c_void_p()
        L.check(L.lib().mrt_scene_create_procedural(name.encode(), int(seed), C.byref(handle)))
        return cls._from_handle(handle, name)

    @classmethod
    def from_npz(cls, path, prefix):
        """Scene stored as ABI arrays under `prefix/` in an .npz (tests/golden/*.npz: the reference's Media files run
        through load_obj where the reference tree is present, so that the GPU box can render them without it)."""
        with np.load(path) as z:
            lights = z[f"{prefix}/lights"]
            return cls(z[f"{prefix}/vertices"], z[f"{prefix}/indices"], z[f"{prefix}/references"], z[f"{prefix}/materials"],
                       lights, len(lights) - 1, f"{prefix} ({path})")

    @classmethod
    def load_obj(cls, path):
        handle = C.c_void_p()
        L.check(L.lib().mrt_scene_load_obj(str(path).encode(), C.byref(handle)))
        return cls._from_handle(handle, str(path))


def noise_tile(seed, frame):
    """64*64 float4 tile for (seed, frame): (mt19937_64(seed ^ frame)() >> 40) * 2^-24."""
    tile = np.empty(L.NOISE_TILE_FLOATS, dtype=np.float32)
    L.lib().mrt_noise_fill(int(seed), int(frame), L.ptr(tile))
    return tile


def noise_tiles(seed, first_frame, count, out=None):
    tiles = out if out is not None else np.empty((count, L.NOISE_TILE_FLOATS), dtype=np.float32)
    for k in range(count):
        L.lib().mrt_noise_fill(int(seed), int(first_frame + k), L.ptr(tiles[k]))
    return tiles


def bvh_selftest(name, seed=0, n_rays=2000, ray_seed=1):
    """Host-only: build the 8-wide BVH of a procedural scene (or an .obj path) and check a host copy of the traversal
    against brute force.  Returns a dict (mismatches must be 0)."""
    handle = C.c_void_p()
    if str(name).endswith(".obj"):
        L.check(L.lib().mrt_scene_load_obj(str(name).encode(), C.byref(handle)))
    else:
        L.check(L.lib().mrt_scene_create_procedural(str(name).encode(), int(seed), C.byref(handle)))
    rep = L.BvhReport()
    try:
        L.check(L.lib().mrt_bvh_selftest(handle, int(n_rays), int(ray_seed), C.byref(rep)))
    finally:
        L.lib().mrt_scene_destroy(handle)
    return {n: getattr(rep, n) for n, _ in L.BvhReport._fields_}


def build_alias_table(pdf):
    """(prob, alias) of the Walker/Vose table mrt_set_scene builds from the light pdfs (optics extension, host code)."""
    pdf = np.ascontiguousarray(pdf, dtype=np.float32)
    prob = np.zeros(len(pdf), dtype=np.float32)
    alias = np.zeros(len(pdf), dtype=np.uint32)
    L.check(L.lib().mrt_build_alias_table(L.ptr(pdf), len(pdf), L.ptr(prob), L.ptr(alias)))
    return prob, alias
