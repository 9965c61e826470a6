"""CPU tests of the acceleration structure the CUDA kernels traverse (no GPU needed): the 8-wide quantised BVH is built
by the product library and a host copy of the traversal state machine — sharing the node decode with the kernel
(csrc/bvh8.h) — must return exactly the brute-force (distance, triangle) for every ray.  Replaces trusting
MPSTriangleAccelerationStructure (reference renderer/Renderer.mm:456-462)."""
import pytest

import mrt_b200 as m


@pytest.mark.parametrize("name,rays", [("cornell", 4000), ("white-box", 2000), ("cornell-water", 2500),
                                       ("spheres-64k", 500), ("many-lights", 600)])
def test_wide_bvh_traversal_equals_brute_force(name, rays):
    rep = m.bvh_selftest(name, seed=3, n_rays=rays, ray_seed=11)
    assert rep["mismatches"] == 0, rep
    assert rep["triangles"] == m.Scene.procedural(name, 3).triangle_count
    assert rep["depth"] < 30 and rep["nodes"] >= 1
    if rep["triangles"] > 1000:
        assert rep["nodes"] * 4 < rep["triangles"]           # 8-wide nodes with up to 3 triangles per leaf
        assert rep["nodes_per_ray"] < 12


def test_wide_bvh_of_reference_media_obj_if_present():
    import os
    p = "/root/reference/renderer/Media/cornellbox.obj"
    if not os.path.exists(p):
        pytest.skip("reference media not on this machine")
    rep = m.bvh_selftest(p, n_rays=3000)
    assert rep["mismatches"] == 0 and rep["triangles"] == 36
