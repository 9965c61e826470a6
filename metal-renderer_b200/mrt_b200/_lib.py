"""ctypes binding of libmrt_b200.so (the C-ABI of include/mrt.h).

There is no CPU fallback: if the shared library is missing this module raises, and if no CUDA device is present
``mrt_create`` fails with ``MRT_E_CUDA``.  Host-only helpers (scene generation, OBJ loading, noise tiles) work
without a GPU.
"""
import ctypes as C
import os
import pathlib

import numpy as np

PKG_DIR = pathlib.Path(__file__).resolve().parent
LIB_PATH = PKG_DIR.parent / "lib" / "libmrt_b200.so"

# numpy views of the ABI structs (include/mrt_abi.h <- reference renderer/Raytracing.h:47-123)
RAY = np.dtype([("origin", "<f4", 3), ("minDistance", "<f4"), ("direction", "<f4", 3), ("maxDistance", "<f4"),
                ("throughput", "<f4", 3), ("radiance", "<f4", 3), ("_pad", "<f4", 2), ("params", "<f4", 4)])
LIGHT_RAY = np.dtype([("origin", "<f4", 3), ("minDistance", "<f4"), ("direction", "<f4", 3), ("maxDistance", "<f4"),
                      ("throughput", "<f4", 3), ("targetIndex", "<u4")])
INTERSECTION = np.dtype([("distance", "<f4"), ("triangleIndex", "<u4"), ("coordinates", "<f4", 2)])
VERTEX = np.dtype([("v", "<f4", 3), ("n", "<f4", 3)])
MATERIAL = np.dtype([("diffuse", "<f4", 3), ("emissive", "<f4", 3), ("ior", "<f4"), ("materialType", "<u4")])
TRIANGLE_REFERENCE = np.dtype([("tri", "<u4", 3), ("materialIndex", "<u4"), ("lightTriangleIndex", "<u4")])
LIGHT_TRIANGLE = np.dtype([("emissive", "<f4", 3), ("v1", VERTEX), ("v2", VERTEX), ("v3", VERTEX),
                           ("area", "<f4"), ("pdf", "<f4"), ("cdf", "<f4"), ("index", "<u4")])
assert (RAY.itemsize, LIGHT_RAY.itemsize, INTERSECTION.itemsize, VERTEX.itemsize, MATERIAL.itemsize,
        TRIANGLE_REFERENCE.itemsize, LIGHT_TRIANGLE.itemsize) == (80, 48, 16, 24, 32, 20, 100)

NOISE_DIMENSIONS = 64
NOISE_TILE_FLOATS = 4 * NOISE_DIMENSIONS * NOISE_DIMENSIONS
NOT_A_LIGHT = 0xFFFFFFFF
MATERIAL_DIFFUSE, MATERIAL_MIRROR, MATERIAL_SMOOTH_PLASTIC, MATERIAL_SMOOTH_DIELECTRIC = range(4)
PIPELINE_AUTO, PIPELINE_MEGAKERNEL, PIPELINE_WAVEFRONT, PIPELINE_STAGED = range(4)
ACCUM_MEAN, ACCUM_SUM, ACCUM_REPLACE = range(3)
COMPARE_DISABLED, COMPARE_ABSOLUTE_VALUE, COMPARE_REF_TO_COLOR, COMPARE_COLOR_TO_REF, COMPARE_LUMINANCE = range(5)   # Raytracing.h:27-31
SCHEDULE_GRAPH, SCHEDULE_HOST_LOOP = 0, 1   # mrt_set_wave_schedule
EXT_REFRACTION, EXT_DISPERSION, EXT_ALIAS_LIGHTS, EXT_CONTINUOUS_SPECTRUM = 1, 2, 4, 8   # mrt_set_extensions (beyond the reference)
NOALLOC_MAX_RAYS, NOALLOC_MAX_STRIDE = 65536, 80   # mrt_intersect_host_noalloc (include/mrt.h)
PROFILE_TRACE_TIME, PROFILE_VISITS = 1, 2      # mrt_set_profile flags (include/mrt.h)
BUF_RAYS, BUF_LIGHT_RAYS, BUF_INTERSECTIONS, BUF_IMAGE = range(4)
BUFFER_DTYPES = {BUF_RAYS: RAY, BUF_LIGHT_RAYS: LIGHT_RAY, BUF_INTERSECTIONS: INTERSECTION}


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("max_path_length", C.c_uint32),
                ("accumulate_mode", C.c_uint32), ("animate_noise", C.c_uint32), ("pipeline", C.c_uint32),
                ("bvh_max_leaf", C.c_uint32), ("bvh_threads", C.c_uint32), ("count_rays", C.c_uint32),
                ("reserved", C.c_uint32 * 7)]


class Stats(C.Structure):
    _fields_ = [("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("frames_rendered", C.c_uint64), ("nodes_visited", C.c_uint64), ("triangles_tested", C.c_uint64),
                ("last_render_ms", C.c_float), ("bvh_nodes", C.c_uint32), ("bvh_depth", C.c_uint32),
                ("bvh_sah_cost", C.c_float), ("bvh_build_ms", C.c_float), ("last_kernel_ms", C.c_float),
                ("bvh8_nodes", C.c_uint32), ("bvh8_depth", C.c_uint32), ("profile_trace_ms", C.c_float),
                ("profile_trace_launches", C.c_uint32), ("bvh8_build_ms", C.c_float), ("waves_traced", C.c_uint32),
                ("tail_waves", C.c_uint32), ("reserved", C.c_uint32)]


class PeerHandle(C.Structure):      # mrt_peer_handle (include/mrt.h)
    _fields_ = [("block_ipc", C.c_ubyte * 64), ("block_ptr", C.c_uint64), ("pid", C.c_int64), ("device", C.c_int32),
                ("width", C.c_uint32), ("height", C.c_uint32), ("ipc_valid", C.c_uint32), ("mapped_everywhere", C.c_uint32),
                ("reserved", C.c_uint32)]


PEER_MAX_RANKS = 16
PEER_SHARE_IPC, PEER_SHARE_PROCESS = 0, 1


class SceneView(C.Structure):
    _fields_ = [("vertices", C.c_void_p), ("vertex_count", C.c_uint32),
                ("indices", C.c_void_p), ("triangle_count", C.c_uint32),
                ("references", C.c_void_p),
                ("materials", C.c_void_p), ("material_count", C.c_uint32),
                ("lights", C.c_void_p), ("light_count", C.c_uint32)]


class BvhReport(C.Structure):
    _fields_ = [("nodes", C.c_uint32), ("depth", C.c_uint32), ("triangles", C.c_uint32), ("mismatches", C.c_uint32),
                ("sah_cost", C.c_float), ("build_ms", C.c_float), ("nodes_per_ray", C.c_float),
                ("triangles_per_ray", C.c_float)]


class DisplayParams(C.Structure):
    """mrt_display_params (include/mrt.h): the presentation switches of Raytracing.h:11-12, 27-33 as parameters."""
    _fields_ = [("tone_mapping", C.c_uint32), ("manual_srgb", C.c_uint32), ("comparison_mode", C.c_uint32),
                ("comparison_scale", C.c_float)]


class MrtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"mrt error {code}: {message}")
        self.code = code


# every symbol include/mrt.h declares: name -> (restype, argtypes)
_P, _U32, _U64, _SZ, _F = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t, C.c_float
SIGNATURES = {
    "mrt_last_error": (C.c_char_p, []),
    "mrt_version": (C.c_int, []),
    "mrt_config_default": (None, [C.POINTER(Config)]),
    "mrt_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "mrt_destroy": (C.c_int, [_P]),
    "mrt_set_scene": (C.c_int, [_P, _P, _U32, _P, _U32, _P, _P, _U32, _P, _U32]),
    "mrt_set_scene_from": (C.c_int, [_P, _P]),
    "mrt_scene_prepare": (C.c_int, [_P, _U32, _P, _U32, _P, _P, _U32, _P, _U32, _U32, _U32, C.POINTER(_P)]),
    "mrt_scene_prepare_from": (C.c_int, [_P, _U32, C.POINTER(_P)]),
    "mrt_set_prepared_scene": (C.c_int, [_P, _P]),
    "mrt_prepared_scene_destroy": (None, [_P]),
    "mrt_resize": (C.c_int, [_P, _U32, _U32]),
    "mrt_set_noise": (C.c_int, [_P, _U32, _P]),
    "mrt_set_row_shard": (C.c_int, [_P, _U32, _U32]),
    "mrt_set_accumulate_mode": (C.c_int, [_P, _U32]),
    "mrt_set_pipeline": (C.c_int, [_P, _U32]),
    "mrt_set_wave_schedule": (C.c_int, [_P, _U32]),
    "mrt_set_extensions": (C.c_int, [_P, _U32, C.c_float]),
    "mrt_peer_export": (C.c_int, [_P, _U32, _P]),
    "mrt_peer_connect": (C.c_int, [_P, _U32, _U32, _P]),
    "mrt_peer_reduce": (C.c_int, [_P, C.c_float]),
    "mrt_peer_result": (C.c_int, [_P, _P, _P]),
    "mrt_peer_status": (C.c_int, [_P]),
    "mrt_peer_read_result": (C.c_int, [_P, _P]),
    "mrt_peer_disconnect": (C.c_int, [_P]),
    "mrt_build_alias_table": (C.c_int, [_P, _U32, _P, _P]),
    "mrt_set_profile": (C.c_int, [_P, _U32]),
    "mrt_raygen": (C.c_int, [_P, _U32]),
    "mrt_intersect": (C.c_int, [_P, _U32]),
    "mrt_shade": (C.c_int, [_P, _U32, _U32]),
    "mrt_shadow_resolve": (C.c_int, [_P]),
    "mrt_accumulate": (C.c_int, [_P, _U32]),
    "mrt_render_frame_staged": (C.c_int, [_P, _U32]),
    "mrt_render_frames": (C.c_int, [_P, _U32, _U32, _U32, _P]),
    "mrt_sync": (C.c_int, [_P]),
    "mrt_event_record": (C.c_int, [_P, _U32]),
    "mrt_event_elapsed_ms": (C.c_int, [_P, _U32, _U32, C.POINTER(C.c_float)]),
    "mrt_read_image": (C.c_int, [_P, _P]),
    "mrt_scale_image": (C.c_int, [_P, _F]),
    "mrt_get_size": (C.c_int, [_P, C.POINTER(_U32), C.POINTER(_U32)]),
    "mrt_display_params_default": (None, [C.POINTER(DisplayParams)]),
    "mrt_display": (C.c_int, [_P, C.POINTER(DisplayParams), _P, _U32, _U32, _P]),
    "mrt_checkpoint_save": (C.c_int, [_P, C.c_char_p]),
    "mrt_checkpoint_load": (C.c_int, [_P, C.c_char_p, C.POINTER(_U32)]),
    "mrt_clear_image": (C.c_int, [_P]),
    "mrt_image_device_ptr": (C.c_int, [_P, C.POINTER(_P), C.POINTER(_SZ)]),
    "mrt_stream_handle": (C.c_int, [_P, C.POINTER(_P)]),
    "mrt_scale_buffer": (C.c_int, [_P, _P, _SZ, _F]),
    "mrt_shard_layout": (C.c_int, [_P, _U32, C.POINTER(_U32), C.POINTER(_SZ)]),
    "mrt_pack_rows": (C.c_int, [_P, _P]),
    "mrt_unpack_rows": (C.c_int, [_P, _P, _U32]),
    "mrt_read_buffer": (C.c_int, [_P, _U32, _P, _SZ]),
    "mrt_write_buffer": (C.c_int, [_P, _U32, _P, _SZ]),
    "mrt_get_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "mrt_reset_stats": (C.c_int, [_P]),
    "mrt_intersect_host": (C.c_int, [_P, _P, _SZ, _U32, _P]),
    "mrt_intersect_host_noalloc": (C.c_int, [_P, _P, _SZ, _U32, _P]),
    "mrt_profile_traversal": (C.c_int, [_P, _P, _SZ, _U32]),
    "mrt_scene_create_procedural": (C.c_int, [C.c_char_p, _U64, C.POINTER(_P)]),
    "mrt_scene_load_obj": (C.c_int, [C.c_char_p, C.POINTER(_P)]),
    "mrt_scene_get": (C.c_int, [_P, C.POINTER(SceneView)]),
    "mrt_scene_destroy": (None, [_P]),
    "mrt_bvh_selftest": (C.c_int, [_P, _U32, _U64, C.POINTER(BvhReport)]),
    "mrt_noise_fill": (None, [_U64, _U32, _P]),
}

_lib = None


def lib():
    """Load libmrt_b200.so (built in-tree by `make -C metal-renderer_b200` / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        path = os.environ.get("MRT_B200_LIB", str(LIB_PATH))
        if not os.path.exists(path):
            raise ImportError(f"{path} not found: build it with `make -C {PKG_DIR.parent}` "
                              "(there is no CPU fallback for the CUDA path)")
        handle = C.CDLL(path)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if a declared symbol is not exported
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise MrtError(rc, lib().mrt_last_error().decode("utf-8", "replace"))


def ptr(array):
    return array.ctypes.data_as(C.c_void_p)
