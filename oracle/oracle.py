"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liboracle*.so.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (metal-renderer_b200/) never imports this.
"""
import ctypes as C
import os
import pathlib
import subprocess

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent

RAY = np.dtype([("origin", "<f4", 3), ("minDistance", "<f4"), ("direction", "<f4", 3), ("maxDistance", "<f4"),
                ("throughput", "<f4", 3), ("radiance", "<f4", 3), ("_pad", "<f4", 2), ("params", "<f4", 4)])
LIGHT_RAY = np.dtype([("origin", "<f4", 3), ("minDistance", "<f4"), ("direction", "<f4", 3), ("maxDistance", "<f4"),
                      ("throughput", "<f4", 3), ("targetIndex", "<u4")])
INTERSECTION = np.dtype([("distance", "<f4"), ("triangleIndex", "<u4"), ("coordinates", "<f4", 2)])
VERTEX = np.dtype([("v", "<f4", 3), ("n", "<f4", 3)])
MATERIAL = np.dtype([("diffuse", "<f4", 3), ("emissive", "<f4", 3), ("ior", "<f4"), ("materialType", "<u4")])
LIGHT_TRIANGLE = np.dtype([("emissive", "<f4", 3), ("v1", VERTEX), ("v2", VERTEX), ("v3", VERTEX),
                           ("area", "<f4"), ("pdf", "<f4"), ("cdf", "<f4"), ("index", "<u4")])
BUF_DTYPES = {0: RAY, 1: LIGHT_RAY, 2: INTERSECTION}
TILE = 4 * 64 * 64


def _has_fma():
    try:
        with open("/proc/cpuinfo") as f:
            flags = next((line for line in f if line.startswith("flags")), "")
        return " fma " in flags + " " and " avx2 " in flags + " "
    except OSError:
        return False


def build():
    subprocess.check_call(["make", "-s", "-C", str(HERE)])


_lib = None


def lib():
    global _lib
    if _lib is None:
        name = "liboracle_fma.so" if _has_fma() and os.environ.get("ORACLE_GENERIC") != "1" else "liboracle.so"
        path = HERE / name
        if not path.exists():
            build()
        h = C.CDLL(str(path))
        P, U, F = C.c_void_p, C.c_uint32, C.c_float
        sig = {
            "orc_create": (P, [U, U, U]), "orc_destroy": (None, [P]), "orc_threads": (U, [P]),
            "orc_set_options": (None, [P, U, U, U]), "orc_set_accumulate_mode": (None, [P, U]),
            "orc_set_row_shard": (None, [P, U, U]),
            "orc_set_scene": (None, [P, P, U, P, U, P, P, U, P, U]), "orc_resize": (None, [P, U, U]),
            "orc_set_noise": (None, [P, U, P]), "orc_raygen": (None, [P, U]), "orc_intersect": (None, [P, U]),
            "orc_shade": (None, [P, U, U]), "orc_shadow_resolve": (None, [P]), "orc_accumulate": (None, [P, U]),
            "orc_render_frame": (None, [P, U]), "orc_render_frames": (C.c_double, [P, U, U, U, P]),
            "orc_read_buffer": (None, [P, U, P, C.c_size_t]), "orc_write_buffer": (None, [P, U, P, C.c_size_t]),
            "orc_read_image": (None, [P, P]), "orc_get_counts": (None, [P, P]), "orc_reset_counts": (None, [P]),
            "orc_intersect_host": (None, [P, P, C.c_size_t, U, P, U]),
            "orc_fresnel": (F, [P, P, F, F]), "orc_balanceHeuristic": (F, [F, F]),
            "orc_triangleSamplePDF": (F, [F, F, F]), "orc_barycentric": (None, [F, F, P]),
            "orc_buildOrthonormalBasis": (None, [P, P, P]), "orc_generateDiffuseBounce": (None, [F, F, P, P]),
            "orc_sincos": (None, [F, P, P]), "orc_reflect": (None, [P, P, P]),
            "orc_interpolate": (None, [P, P, P, F, F, P]), "orc_selectLightTriangle": (U, [F, P, C.c_int]),
            "orc_sampleMaterial": (None, [P, P, P, P, P, P]),
            "orc_generateNextBounce": (None, [P, P, F, P, P, P, P]),
            "orc_lightTriangleSamplePDF": (F, [F, F, P, P, P]),
            "orc_triangle_test": (None, [P, P, F, F, P, P, P, P, P]), "orc_sizeof": (U, [U]),
            "orc_display": (None, [P, U, U, U, U, U, F, P, U, U, P]), "orc_toSRGB": (F, [F]),
            "orc_set_extensions": (None, [P, U, F, P, P, U]),
            "orc_refractiveBounce": (None, [P, U, F, P, F, P, P, P, P, P, P]),
            "orc_selectLightAlias": (U, [F, P, P, U]),
        }
        for n, (r, a) in sig.items():
            fn = getattr(h, n)
            fn.restype, fn.argtypes = r, a
        h._name = name
        _lib = h
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def f32(*v):
    return np.array(v, dtype=np.float32)


class Oracle:
    """CPU restatement of the reference pipeline; same lifecycle as the product's Renderer."""

    def __init__(self, max_path_length=8, accumulate_mode=0, threads=0):
        self._c = lib().orc_create(max_path_length, accumulate_mode, threads)
        self.width = self.height = 0

    def close(self):
        if getattr(self, "_c", None):
            lib().orc_destroy(self._c)
            self._c = None

    __del__ = close

    @property
    def threads(self):
        return lib().orc_threads(self._c)

    def set_options(self, use_bvh=True, count_visits=False, animate_noise=True):
        lib().orc_set_options(self._c, int(use_bvh), int(count_visits), int(animate_noise))

    def set_accumulate_mode(self, mode):
        lib().orc_set_accumulate_mode(self._c, mode)

    def set_row_shard(self, first, stride):
        lib().orc_set_row_shard(self._c, first, stride)

    def set_extensions(self, flags, dispersion=0.0, alias_prob=None, alias_index=None):
        """Optics extension (beyond the reference, orc_shading.h).  The alias table comes from the caller."""
        if alias_prob is None:
            lib().orc_set_extensions(self._c, flags, dispersion, None, None, 0)
        else:
            prob = np.ascontiguousarray(alias_prob, dtype=np.float32)
            idx = np.ascontiguousarray(alias_index, dtype=np.uint32)
            lib().orc_set_extensions(self._c, flags, dispersion, _p(prob), _p(idx), len(prob))

    def init_raytracing(self, scene):
        self._scene = scene     # keep arrays alive during the call only; oracle copies them
        lib().orc_set_scene(self._c, _p(scene.vertices), len(scene.vertices), _p(scene.indices), len(scene.references),
                            _p(scene.references), _p(scene.materials), len(scene.materials), _p(scene.lights),
                            scene.light_count)

    def drawable_size_will_change(self, w, h):
        lib().orc_resize(self._c, w, h)
        self.width, self.height = w, h

    def update_noise(self, slot, tile):
        tile = np.ascontiguousarray(tile, dtype=np.float32)
        lib().orc_set_noise(self._c, slot, _p(tile))

    def raygen(self, frame): lib().orc_raygen(self._c, frame)
    def intersect(self, which=0): lib().orc_intersect(self._c, which)
    def shade(self, frame, it): lib().orc_shade(self._c, frame, it)
    def shadow_resolve(self): lib().orc_shadow_resolve(self._c)
    def accumulate(self, frame): lib().orc_accumulate(self._c, frame)
    def perform_raytracing(self, frame): lib().orc_render_frame(self._c, frame)

    def draw(self, n_frames, noise_tiles=None, first_frame=0, frame_stride=1):
        p = None
        if noise_tiles is not None:
            noise_tiles = np.ascontiguousarray(noise_tiles, dtype=np.float32)
            p = _p(noise_tiles)
        return lib().orc_render_frames(self._c, first_frame, n_frames, frame_stride, p)

    def read_buffer(self, which):
        n = self.width * self.height
        out = np.empty((n, 4), np.float32) if which == 3 else np.empty(n, BUF_DTYPES[which])
        lib().orc_read_buffer(self._c, which, _p(out), out.nbytes)
        return out

    def write_buffer(self, which, arr):
        arr = np.ascontiguousarray(arr)
        lib().orc_write_buffer(self._c, which, _p(arr), arr.nbytes)

    def save_current_image(self):
        img = np.empty((self.height, self.width, 4), np.float32)
        lib().orc_read_image(self._c, _p(img))
        return img

    def counts(self):
        out = np.zeros(4, np.uint64)
        lib().orc_get_counts(self._c, _p(out))
        return dict(primary_rays=int(out[0]), shadow_rays=int(out[1]), nodes_visited=int(out[2]), triangles_tested=int(out[3]))

    def reset_counts(self):
        lib().orc_reset_counts(self._c)

    def intersect_host(self, rays, stride=None, bvh=True):
        rays = np.ascontiguousarray(rays)
        stride = stride or rays.dtype.itemsize
        n = rays.nbytes // stride
        out = np.empty(n, INTERSECTION)
        lib().orc_intersect_host(self._c, _p(rays), stride, n, _p(out), int(bvh))
        return out


def display(image, tone_mapping=False, manual_srgb=False, comparison_mode=0, comparison_scale=10.0, reference=None):
    """Presentation pass of the reference (Shaders.metal:33-69) on an (H, W, 4) float32 image; reference = (RH, RW, 4)."""
    image = np.ascontiguousarray(image, np.float32)
    h, w = image.shape[:2]
    out = np.empty_like(image)
    if reference is not None:
        reference = np.ascontiguousarray(reference, np.float32)
        rh, rw = reference.shape[:2]
        lib().orc_display(_p(image), w, h, int(tone_mapping), int(manual_srgb), int(comparison_mode), float(comparison_scale),
                          _p(reference), rw, rh, _p(out))
    else:
        lib().orc_display(_p(image), w, h, int(tone_mapping), int(manual_srgb), int(comparison_mode), float(comparison_scale),
                          None, 0, 0, _p(out))
    return out
