"""mrt_b200 — host-side Python mirror of the reference's Renderer interface over libmrt_b200.so."""
from ._lib import (EXT_ALIAS_LIGHTS, EXT_CONTINUOUS_SPECTRUM, EXT_DISPERSION, EXT_REFRACTION, ACCUM_MEAN, ACCUM_REPLACE, ACCUM_SUM, BUF_IMAGE, COMPARE_ABSOLUTE_VALUE, COMPARE_COLOR_TO_REF,
                   COMPARE_DISABLED, COMPARE_LUMINANCE, COMPARE_REF_TO_COLOR, BUF_INTERSECTIONS, BUF_LIGHT_RAYS, BUF_RAYS,
                   INTERSECTION, LIGHT_RAY, LIGHT_TRIANGLE, MATERIAL, NOISE_TILE_FLOATS, NOT_A_LIGHT, PIPELINE_AUTO,
                   PIPELINE_MEGAKERNEL, PIPELINE_STAGED, PIPELINE_WAVEFRONT, PROFILE_TRACE_TIME, PROFILE_VISITS, RAY, SCHEDULE_GRAPH, SCHEDULE_HOST_LOOP, TRIANGLE_REFERENCE, VERTEX, MrtError, lib)
from .renderer import Renderer
from .scene import PROCEDURAL_SCENES, Scene, build_alias_table, bvh_selftest, noise_tile, noise_tiles

__all__ = [n for n in dir() if not n.startswith("_")]
