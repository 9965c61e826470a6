// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_abi.h).
//
// orc_bvh.h — nearest-hit ray/triangle intersector standing in for Apple's closed-source
// MPSRayIntersector / MPSTriangleAccelerationStructure (call sites: renderer/Renderer.mm:456-469,
// 519-523, 545-553).  PARITY UNPINNED against MPS itself (no source, no vectors); what IS pinned:
//   * ray layout OriginMinDistanceDirectionMaxDistance read with a byte stride (Renderer.mm:466-467),
//   * output DistancePrimitiveIndexCoordinates (Renderer.mm:469), coordinates = weights of vertex 0 and 1
//     (consumed by KernelHelpers.h:39-42: P = u*V0 + v*V1 + (1-u-v)*V2),
//   * no culling (Renderer.mm:465), no masks (:468), rays with maxDistance < 0 skipped, miss => distance < 0.
// The ray/triangle test is Moeller-Trumbore anchored at V2 in the pinned arithmetic of orc_math.h; the
// nearest hit is min over ALL triangles of (t, triangleIndex) lexicographically, so the result does not
// depend on any acceleration structure.  Two evaluators are provided: brute force (the definition) and an
// oracle-private binned-SAH BVH (for scenes too big for brute force); tests check they agree bit-for-bit.
#pragma once
#include <algorithm>
#include <cfloat>
#include <vector>
#include "orc_abi.h"
#include "orc_math.h"

namespace orc {

struct Hit { float t; uint32_t tri; float u, v; };

// One ray/triangle test.  V0,V1,V2 = vertex positions of triangle `tri` (vertex buffer stride 24 +
// uint32 indices, Renderer.mm:457-461).  Updates `best` iff the triangle is hit inside [tmin, tmax] and
// (t, tri) < (best.t, best.tri).
static inline void triangle_test(f3 o, f3 d, float tmin, float tmax, f3 V0, f3 V1, f3 V2, uint32_t tri, Hit& best) {
    f3 e1 = V0 - V2;
    f3 e2 = V1 - V2;
    f3 p = cross(d, e2);
    float det = dot(e1, p);
    float inv = 1.0f / det;                 // det == 0 -> inf/NaN, rejected by the range tests below
    f3 s = o - V2;
    float u = dot(s, p) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return;
    f3 q = cross(s, e1);
    float v = dot(d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return;
    float t = dot(e2, q) * inv;
    if (!(t >= tmin && t <= tmax)) return;
    if (t < best.t || (t == best.t && tri < best.tri)) best = Hit{t, tri, u, v};
}

struct Bvh {
    struct Node { float lo[3], hi[3]; uint32_t left, right, first, count; };   // count>0 => leaf
    std::vector<Node> nodes;
    std::vector<uint32_t> order;          // triangle indices in leaf order
    const Vertex* verts = nullptr;
    const uint32_t* idx = nullptr;
    uint32_t ntri = 0;
    float pad = 0.0f;

    f3 P(uint32_t tri, int k) const { return ld3(verts[idx[3 * tri + k]].v); }

    void build(const Vertex* v, const uint32_t* i, uint32_t n) {
        verts = v; idx = i; ntri = n;
        nodes.clear(); order.resize(n);
        if (n == 0) return;
        std::vector<float> cen(3 * (size_t)n), blo(3 * (size_t)n), bhi(3 * (size_t)n);
        float slo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, shi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (uint32_t t = 0; t < n; ++t) {
            order[t] = t;
            for (int a = 0; a < 3; ++a) {
                float p0 = verts[idx[3 * t]].v[a], p1 = verts[idx[3 * t + 1]].v[a], p2 = verts[idx[3 * t + 2]].v[a];
                float lo = std::min(p0, std::min(p1, p2)), hi = std::max(p0, std::max(p1, p2));
                blo[3 * (size_t)t + a] = lo; bhi[3 * (size_t)t + a] = hi; cen[3 * (size_t)t + a] = 0.5f * (lo + hi);
                slo[a] = std::min(slo[a], lo); shi[a] = std::max(shi[a], hi);
            }
        }
        // Conservative padding: a float "hit" may be a mathematical near-miss by a few ulps; node boxes are
        // inflated so such triangles are still reached (keeps the BVH answer == brute force).
        float ext = std::max(shi[0] - slo[0], std::max(shi[1] - slo[1], shi[2] - slo[2]));
        float mag = 0.0f;
        for (int a = 0; a < 3; ++a) mag = std::max(mag, std::max(std::fabs(slo[a]), std::fabs(shi[a])));
        pad = 1e-5f * std::max(ext, mag) + 1e-30f;
        nodes.reserve(2 * (size_t)n);
        nodes.push_back(Node{});
        subdivide(0, 0, n, cen, blo, bhi);
    }

    void bounds(uint32_t first, uint32_t count, const std::vector<float>& blo, const std::vector<float>& bhi,
                float* lo, float* hi) const {
        for (int a = 0; a < 3; ++a) { lo[a] = FLT_MAX; hi[a] = -FLT_MAX; }
        for (uint32_t k = first; k < first + count; ++k)
            for (int a = 0; a < 3; ++a) {
                lo[a] = std::min(lo[a], blo[3 * (size_t)order[k] + a]);
                hi[a] = std::max(hi[a], bhi[3 * (size_t)order[k] + a]);
            }
    }

    void subdivide(uint32_t ni, uint32_t first, uint32_t count, const std::vector<float>& cen,
                   const std::vector<float>& blo, const std::vector<float>& bhi) {
        float lo[3], hi[3];
        bounds(first, count, blo, bhi, lo, hi);
        for (int a = 0; a < 3; ++a) { nodes[ni].lo[a] = lo[a] - pad; nodes[ni].hi[a] = hi[a] + pad; }
        nodes[ni].first = first; nodes[ni].count = count; nodes[ni].left = nodes[ni].right = 0;
        if (count <= 4) return;
        // binned SAH over centroid bounds, 12 bins, all three axes
        float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (uint32_t k = first; k < first + count; ++k)
            for (int a = 0; a < 3; ++a) {
                float c = cen[3 * (size_t)order[k] + a];
                clo[a] = std::min(clo[a], c); chi[a] = std::max(chi[a], c);
            }
        const int NB = 12;
        int bestAxis = -1, bestBin = 0; float bestCost = FLT_MAX;
        for (int a = 0; a < 3; ++a) {
            if (!(chi[a] > clo[a])) continue;
            float scale = NB / (chi[a] - clo[a]);
            float bl[NB][3], bh[NB][3]; uint32_t bc[NB];
            for (int b = 0; b < NB; ++b) { bc[b] = 0; for (int k = 0; k < 3; ++k) { bl[b][k] = FLT_MAX; bh[b][k] = -FLT_MAX; } }
            for (uint32_t k = first; k < first + count; ++k) {
                size_t t = order[k];
                int b = std::min(NB - 1, (int)((cen[3 * t + a] - clo[a]) * scale));
                bc[b]++;
                for (int c = 0; c < 3; ++c) { bl[b][c] = std::min(bl[b][c], blo[3 * t + c]); bh[b][c] = std::max(bh[b][c], bhi[3 * t + c]); }
            }
            float la[NB], ra[NB]; uint32_t lc[NB], rc[NB];
            float l[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, h[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}; uint32_t c = 0;
            auto area = [](const float* l, const float* h) {
                float dx = h[0] - l[0], dy = h[1] - l[1], dz = h[2] - l[2];
                return dx * dy + dy * dz + dz * dx;
            };
            for (int b = 0; b < NB - 1; ++b) {
                for (int k = 0; k < 3; ++k) { l[k] = std::min(l[k], bl[b][k]); h[k] = std::max(h[k], bh[b][k]); }
                c += bc[b]; lc[b] = c; la[b] = c ? area(l, h) : 0.0f;
            }
            for (int k = 0; k < 3; ++k) { l[k] = FLT_MAX; h[k] = -FLT_MAX; }
            c = 0;
            for (int b = NB - 1; b > 0; --b) {
                for (int k = 0; k < 3; ++k) { l[k] = std::min(l[k], bl[b][k]); h[k] = std::max(h[k], bh[b][k]); }
                c += bc[b]; rc[b - 1] = c; ra[b - 1] = c ? area(l, h) : 0.0f;
            }
            for (int b = 0; b < NB - 1; ++b) {
                if (!lc[b] || !rc[b]) continue;
                float cost = la[b] * lc[b] + ra[b] * rc[b];
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = b; }
            }
        }
        uint32_t mid;
        if (bestAxis < 0) {
            mid = first + count / 2;     // all centroids coincide: split by position in the list
        } else {
            float scale = NB / (chi[bestAxis] - clo[bestAxis]);
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t t) {
                int b = std::min(NB - 1, (int)((cen[3 * (size_t)t + bestAxis] - clo[bestAxis]) * scale));
                return b <= bestBin;
            });
            mid = (uint32_t)(it - order.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        uint32_t L = (uint32_t)nodes.size(); nodes.push_back(Node{});
        uint32_t R = (uint32_t)nodes.size(); nodes.push_back(Node{});
        nodes[ni].left = L; nodes[ni].right = R; nodes[ni].count = 0;
        subdivide(L, first, mid - first, cen, blo, bhi);
        subdivide(R, mid, first + count - mid, cen, blo, bhi);
    }

    // slab test; NaNs (0*inf) drop out of min/max so the axis is ignored (conservative)
    static inline bool box_hit(const Node& n, f3 o, f3 inv, float tmin, float tmax, float& tnear) {
        float t0 = tmin, t1 = tmax;
        const float oo[3] = {o.x, o.y, o.z}, ii[3] = {inv.x, inv.y, inv.z};
        for (int a = 0; a < 3; ++a) {
            float ta = (n.lo[a] - oo[a]) * ii[a], tb = (n.hi[a] - oo[a]) * ii[a];
            float tn = ta < tb ? ta : tb, tf = ta < tb ? tb : ta;
            if (ta != ta || tb != tb) continue;
            if (tn > t0) t0 = tn;
            if (tf < t1) t1 = tf;
        }
        tnear = t0;
        return t0 <= t1 * 1.0000004f;
    }

    Hit nearest(f3 o, f3 d, float tmin, float tmax, uint64_t* nodesVisited, uint64_t* trisTested) const {
        const Hit miss{-1.0f, 0xFFFFFFFFu, 0.0f, 0.0f};
        Hit best{tmax, 0xFFFFFFFFu, 0.0f, 0.0f};
        if (nodes.empty()) return miss;
        f3 inv{1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
        struct Item { uint32_t node; float tnear; };
        Item stack[128]; int sp = 0;
        float tn;
        if (nodesVisited) ++*nodesVisited;
        if (box_hit(nodes[0], o, inv, tmin, best.t, tn)) stack[sp++] = Item{0, tn};
        while (sp) {
            Item it = stack[--sp];
            if (it.tnear > best.t) continue;         // non-strict keep: equal-t ties must still be tested
            const Node& n = nodes[it.node];
            if (n.count) {
                for (uint32_t k = n.first; k < n.first + n.count; ++k) {
                    uint32_t t = order[k];
                    if (trisTested) ++*trisTested;
                    triangle_test(o, d, tmin, tmax, P(t, 0), P(t, 1), P(t, 2), t, best);
                }
            } else {
                float tl, tr;
                if (nodesVisited) *nodesVisited += 2;
                bool hl = box_hit(nodes[n.left], o, inv, tmin, best.t, tl);
                bool hr = box_hit(nodes[n.right], o, inv, tmin, best.t, tr);
                if (hl && hr) {
                    if (tl <= tr) { stack[sp++] = Item{n.right, tr}; stack[sp++] = Item{n.left, tl}; }
                    else          { stack[sp++] = Item{n.left, tl}; stack[sp++] = Item{n.right, tr}; }
                } else if (hl) stack[sp++] = Item{n.left, tl};
                else if (hr) stack[sp++] = Item{n.right, tr};
            }
        }
        return best.tri == 0xFFFFFFFFu ? miss : best;
    }

    Hit nearest_brute(f3 o, f3 d, float tmin, float tmax) const {
        Hit best{tmax, 0xFFFFFFFFu, 0.0f, 0.0f};
        for (uint32_t t = 0; t < ntri; ++t) triangle_test(o, d, tmin, tmax, P(t, 0), P(t, 1), P(t, 2), t, best);
        if (best.tri == 0xFFFFFFFFu) return Hit{-1.0f, 0xFFFFFFFFu, 0.0f, 0.0f};
        return best;
    }
};

}  // namespace orc
