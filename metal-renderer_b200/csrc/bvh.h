// bvh.h — host SAH BVH builder producing the flat device layout the traversal kernels read.
// Replaces MPSTriangleAccelerationStructure -rebuild (reference: renderer/Renderer.mm:456-462; closed source).
#pragma once
#include <cstdint>
#include <vector>
#include "mrt_abi.h"

namespace mrt {

// One 64-byte node = both children's boxes + both child links (4 x 128-bit loads).
//   q0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
//   q1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//   q2 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
//   q3 = (child0, child1, 0, 0) as int32: >= 0 inner node index; < 0 leaf, ~code = (firstTri << 3) | count
struct BvhNode { float q[12]; int32_t child[2]; int32_t pad[2]; };
static_assert(sizeof(BvhNode) == 64, "node is 64 bytes");

// One 48-byte triangle record, in leaf order (3 x 128-bit loads):
//   (V2.xyz, bits(original triangle index)), (V0-V2, 0), (V1-V2, 0)   [Moeller-Trumbore anchored at V2]
struct BvhTri { float v2[3]; uint32_t index; float e1[3]; float pad1; float e2[3]; float pad2; };
static_assert(sizeof(BvhTri) == 48, "triangle record is 48 bytes");

struct BvhResult {
    std::vector<BvhNode> nodes;     // nodes[0] = root
    std::vector<BvhTri> tris;
    uint32_t depth = 0;
    float sah_cost = 0.0f;          // sum over inner nodes of area/rootArea + sum over leaves of count*area/rootArea
    float pad = 0.0f;               // conservative box inflation that was applied
    double build_ms = 0.0;
};

// vertices: stride sizeof(mrt_vertex) = 24 (setVertexStride, Renderer.mm:459); indices: uint32 (:461).
void build_bvh(const mrt_vertex* vertices, const uint32_t* indices, uint32_t triangle_count,
               uint32_t max_leaf, uint32_t threads, BvhResult& out);

}  // namespace mrt
