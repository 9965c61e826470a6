// device_trace.cuh — nearest-hit BVH traversal + ray/triangle test (sm_100a device code).
//
// Replaces MPSRayIntersector -encodeIntersection... intersectionType Nearest (reference call sites:
// renderer/Renderer.mm:464-469 config, 519-523 primary rays, 545-553 light-sampling rays; implementation closed).
// Semantics kept: ray = origin, minDistance, direction, maxDistance; no culling; rays with maxDistance < 0 are
// skipped by the caller; result = distance, primitive index, barycentric weights of vertices 0 and 1; miss =>
// distance < 0.  The answer is defined as min over all triangles of (t, triangle index) under the
// Moeller-Trumbore test below, so it does not depend on the tree (boxes are inflated, pruning is non-strict).
//
// Layout: 64-byte nodes holding both children's boxes (4 x LDG.128) and 48-byte triangle records
// (3 x LDG.128), see bvh.h.  Per-thread traversal stack in local memory (L1-resident).
#pragma once
#include "device_shading.cuh"

namespace mrt {

#define MRT_STACK_DEPTH 64
#define MRT_NO_HIT 0xFFFFFFFFu

// One ray/triangle test against leaf-ordered record k.  Updates `best` iff hit within [tmin, tmax] and
// (t, index) < (best.t, best.tri).
__device__ __forceinline__ void triangle_test(const float4* __restrict__ tris, uint32_t k, f3 o, f3 d, float tmin, float tmax,
                                              HitRec& best) {
    const float4* p = tris + 3 * (size_t)k;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    f3 V2 = mk3(a.x, a.y, a.z), e1 = mk3(b.x, b.y, b.z), e2 = mk3(c.x, c.y, c.z);
    uint32_t index = __float_as_uint(a.w);
    f3 pv = cross(d, e2);
    float det = dot(e1, pv);
    float inv = __fdiv_rn(1.0f, det);            // det == 0 -> inf/NaN, rejected by the range tests
    f3 s = o - V2;
    float u = dot(s, pv) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return;
    f3 q = cross(s, e1);
    float v = dot(d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return;
    float t = dot(e2, q) * inv;
    if (!(t >= tmin && t <= tmax)) return;
    if (t < best.t || (t == best.t && index < best.tri)) { best.t = t; best.tri = index; best.u = u; best.v = v; }
}

// Nearest hit of one ray.  COUNT adds node / triangle visit counters (profiling variant only).
template <bool COUNT>
__device__ __forceinline__ HitRec trace_nearest(const float4* __restrict__ nodes, const float4* __restrict__ tris,
                                                f3 o, f3 d, float tmin, float tmax,
                                                unsigned long long* nodesVisited, unsigned long long* trisTested) {
    HitRec best; best.t = tmax; best.tri = MRT_NO_HIT; best.u = 0.0f; best.v = 0.0f;
    // slab test in fma form: t = plane * (1/d) - o * (1/d).  Not bit-matched with anything — only conservative
    // (the boxes carry the builder's padding).  Direction components are kept away from zero FOR THE BOX TEST
    // ONLY, so 1/d stays finite: with an infinite reciprocal plane*inf - o*inf is NaN for one of the two planes
    // and fminf/fmaxf would then turn an axis-parallel ray inside the slab into a miss.
    const float tiny = 1e-30f;
    const float idx = __fdiv_rn(1.0f, fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x);
    const float idy = __fdiv_rn(1.0f, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y);
    const float idz = __fdiv_rn(1.0f, fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
    const float ox = o.x * idx, oy = o.y * idy, oz = o.z * idz;
    int stack[MRT_STACK_DEPTH];
    int sp = 0;
    int cur = 0;
    for (;;) {
        while (cur >= 0) {
            const float4* np = nodes + 4 * (size_t)cur;
            float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
            if (COUNT) ++*nodesVisited;
            float c0x0 = __fmaf_rn(n0.x, idx, -ox), c0x1 = __fmaf_rn(n0.y, idx, -ox);
            float c0y0 = __fmaf_rn(n0.z, idy, -oy), c0y1 = __fmaf_rn(n0.w, idy, -oy);
            float c0z0 = __fmaf_rn(n2.x, idz, -oz), c0z1 = __fmaf_rn(n2.y, idz, -oz);
            float c1x0 = __fmaf_rn(n1.x, idx, -ox), c1x1 = __fmaf_rn(n1.y, idx, -ox);
            float c1y0 = __fmaf_rn(n1.z, idy, -oy), c1y1 = __fmaf_rn(n1.w, idy, -oy);
            float c1z0 = __fmaf_rn(n2.z, idz, -oz), c1z1 = __fmaf_rn(n2.w, idz, -oz);
            float tn0 = fmaxf(fmaxf(fminf(c0x0, c0x1), fminf(c0y0, c0y1)), fmaxf(fminf(c0z0, c0z1), tmin));
            float tf0 = fminf(fminf(fmaxf(c0x0, c0x1), fmaxf(c0y0, c0y1)), fminf(fmaxf(c0z0, c0z1), best.t));
            float tn1 = fmaxf(fmaxf(fminf(c1x0, c1x1), fminf(c1y0, c1y1)), fmaxf(fminf(c1z0, c1z1), tmin));
            float tf1 = fminf(fminf(fmaxf(c1x0, c1x1), fmaxf(c1y0, c1y1)), fminf(fmaxf(c1z0, c1z1), best.t));
            bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;       // non-strict: equal-t ties must still be tested
            int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                if (tn1 < tn0) { int t = c0; c0 = c1; c1 = t; }
                stack[sp++] = c1;
                cur = c0;
            } else if (h0 || h1) {
                cur = h0 ? c0 : c1;
            } else {
                if (sp == 0) goto done;
                cur = stack[--sp];
            }
        }
        {
            uint32_t code = ~uint32_t(cur);
            uint32_t first = code >> 3, count = code & 7u;
            for (uint32_t k = 0; k < count; ++k) {
                if (COUNT) ++*trisTested;
                triangle_test(tris, first + k, o, d, tmin, tmax, best);
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
done:
    if (best.tri == MRT_NO_HIT) best.t = -1.0f;
    return best;
}

// ---- persistent warp-level traversal with dynamic ray fetch ---------------------------------------------------
// The per-ray routine above makes a warp wait for its slowest ray (ncu on the 1M-triangle scene: 5.4 of 32 lanes
// active inside the node loop).  Here a warp owns 32 traversal contexts and keeps pulling work items from a
// global counter: whenever fewer than `refillBelow` lanes still hold a ray, the idle lanes fetch new ones, so
// the node/leaf loop runs with most lanes busy regardless of how ray lengths are distributed.  Per ray the
// arithmetic — and therefore the answer, min over all triangles of (t, index) — is that of trace_nearest.
//
// IO policy:
//   bool load(uint32_t item, TraceJob& job)   fetch work item; false = nothing to traverse (policy already
//                                             stored the result, e.g. a skipped ray)
//   void store(uint32_t tag, uint32_t target, const HitRec& best)      target == MRT_NO_HIT for nearest-hit jobs
// A job is either NEAREST (full record) or ANYHIT: best starts as (tTarget, target) and the ray ends at the first
// triangle that beats it lexicographically, i.e. exactly when the nearest hit would not be the target.
struct TraceJob {
    f3 o, d;
    float tmin, tmax;
    uint32_t target;        // MRT_NO_HIT for nearest-hit jobs
    uint32_t tag;           // handed back to store() (ray index / path slot)
};
#define MRT_TRACE_DONE ((int)0x80000000)

template <class IO>
__device__ __forceinline__ void trace_persistent(const float4* __restrict__ nodes, const float4* __restrict__ tris,
                                                 uint32_t total, unsigned int* head, uint32_t refillBelow, uint32_t leafMin, IO& io) {
    const uint32_t lane = threadIdx.x & 31u;
    leafMin &= 0xffu;                                  // upper bits tune the wide-tree kernel only (device_trace8.cuh)
    const float tiny = 1e-30f;
    bool alive = false, exhausted = (total == 0);
    uint32_t tag = 0, target = MRT_NO_HIT;
    f3 o = mk3(0.f, 0.f, 0.f), d = o;
    float idx = 0.f, idy = 0.f, idz = 0.f, ox = 0.f, oy = 0.f, oz = 0.f, tmin = 0.f, tmax = 0.f;
    HitRec best; best.t = 0.f; best.tri = MRT_NO_HIT; best.u = best.v = 0.f;
    int stack[MRT_STACK_DEPTH];
    int sp = 0, cur = MRT_TRACE_DONE;

    for (;;) {
        // ---- refill idle lanes until enough of the warp is busy (or the queue is empty) ----
        unsigned am;
        for (;;) {
            am = __ballot_sync(0xffffffffu, alive);
            if (exhausted || __popc(am) >= (int)refillBelow) break;
            unsigned need = ~am;
            uint32_t n = __popc(need), base = 0;
            int leader = __ffs(need) - 1;
            if ((int)lane == leader) base = atomicAdd(head, n);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!alive) {
                uint32_t it = base + __popc(need & ((1u << lane) - 1u));
                if (it < total) {
                    TraceJob job;
                    if (io.load(it, job)) {
                        tag = job.tag; o = job.o; d = job.d; tmin = job.tmin; tmax = job.tmax; target = job.target;
                        idx = __fdiv_rn(1.0f, fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x);
                        idy = __fdiv_rn(1.0f, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y);
                        idz = __fdiv_rn(1.0f, fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
                        ox = o.x * idx; oy = o.y * idy; oz = o.z * idz;
                        best.t = tmax; best.tri = target; best.u = 0.0f; best.v = 0.0f;
                        stack[0] = MRT_TRACE_DONE; sp = 1; cur = 0;
                        alive = true;
                    }
                }
            }
            exhausted = base + n >= total;
        }
        if (am == 0) break;                               // queue empty and every lane idle

        // ---- traverse.  Every round the warp votes: lanes standing on an inner node want a NODE step, lanes
        // standing on a leaf (or on the end marker) want a LEAF step.  Node steps outnumber leaf steps ~10:1,
        // so the warp keeps taking node steps — with every lane that wants one — until at least `leafMin` lanes
        // are parked on leaves (or nobody wants a node); only then is the leaf code run.  A per-lane
        // "descend until the next leaf" loop instead makes the warp wait for its slowest descender each round
        // (ncu: 8.3 of 32 lanes active in the node loop). ----
        for (;;) {
            const bool wantNode = alive && cur >= 0;
            const unsigned mN = __ballot_sync(0xffffffffu, wantNode);
            const unsigned mL = __ballot_sync(0xffffffffu, alive && cur < 0);
            if ((mN | mL) == 0) break;
            if (mN != 0 && __popc(mL) < (int)leafMin) {
                if (wantNode) {
                    const float4* np = nodes + 4 * (size_t)cur;
                    float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
                    float c0x0 = __fmaf_rn(n0.x, idx, -ox), c0x1 = __fmaf_rn(n0.y, idx, -ox);
                    float c0y0 = __fmaf_rn(n0.z, idy, -oy), c0y1 = __fmaf_rn(n0.w, idy, -oy);
                    float c0z0 = __fmaf_rn(n2.x, idz, -oz), c0z1 = __fmaf_rn(n2.y, idz, -oz);
                    float c1x0 = __fmaf_rn(n1.x, idx, -ox), c1x1 = __fmaf_rn(n1.y, idx, -ox);
                    float c1y0 = __fmaf_rn(n1.z, idy, -oy), c1y1 = __fmaf_rn(n1.w, idy, -oy);
                    float c1z0 = __fmaf_rn(n2.z, idz, -oz), c1z1 = __fmaf_rn(n2.w, idz, -oz);
                    float tn0 = fmaxf(fmaxf(fminf(c0x0, c0x1), fminf(c0y0, c0y1)), fmaxf(fminf(c0z0, c0z1), tmin));
                    float tf0 = fminf(fminf(fmaxf(c0x0, c0x1), fmaxf(c0y0, c0y1)), fminf(fmaxf(c0z0, c0z1), best.t));
                    float tn1 = fmaxf(fmaxf(fminf(c1x0, c1x1), fminf(c1y0, c1y1)), fmaxf(fminf(c1z0, c1z1), tmin));
                    float tf1 = fminf(fminf(fmaxf(c1x0, c1x1), fmaxf(c1y0, c1y1)), fminf(fmaxf(c1z0, c1z1), best.t));
                    bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;       // non-strict: equal-t ties must still be tested
                    int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
                    if (h0 && h1) {
                        if (tn1 < tn0) { int t = c0; c0 = c1; c1 = t; }
                        stack[sp++] = c1;
                        cur = c0;
                    } else if (h0 || h1) {
                        cur = h0 ? c0 : c1;
                    } else {
                        cur = stack[--sp];
                    }
                }
                continue;
            }
            if (alive && cur < 0) {
                if (cur != MRT_TRACE_DONE) {
                    uint32_t code = ~uint32_t(cur);
                    uint32_t first = code >> 3, count = code & 7u;
                    for (uint32_t k = 0; k < count; ++k) triangle_test(tris, first + k, o, d, tmin, tmax, best);
                    // an any-hit job is decided as soon as something beats (tTarget, target)
                    cur = (target != MRT_NO_HIT && best.tri != target) ? MRT_TRACE_DONE : stack[--sp];
                }
                if (cur == MRT_TRACE_DONE) {
                    if (best.tri == MRT_NO_HIT) best.t = -1.0f;
                    io.store(tag, target, best);
                    alive = false;
                }
            }
            unsigned a2 = __ballot_sync(0xffffffffu, alive);
            if (a2 == 0 || (!exhausted && __popc(a2) < (int)refillBelow)) break;
        }
    }
}

}  // namespace mrt
