// device_trace8.cuh — persistent, vote-scheduled traversal of the 8-wide quantised BVH (bvh8.h) on sm_100a.
//
// Stands in for MPSRayIntersector nearest-hit (reference call sites renderer/Renderer.mm:464-469, 519-523, 545-553).
// Same contract as trace_persistent in device_trace.cuh (IO policy, NEAREST / ANYHIT jobs, answer = min over all
// triangles of (t, index) under triangle_test), different tree: one node fetch (5 x LDG.128) tests 8 children, so
// the chain of dependent loads per ray is ~3x shorter than in the binary tree — the stall that dominated it.
//
// Per-lane state: node group G = (childBase, hits<<24 | imask) of inner children still to visit, triangle group
// T = (triBase, 24-bit mask) of leaf triangles still to test, and a stack of node groups in local memory.
// Every round the warp votes between a NODE step (pop the nearest child of G, fetch it, build new G/T) and a
// TRIANGLE step (drain T); lanes that finish are refilled from the global work counter.
#pragma once
#include "bvh8.h"
#include "device_trace.cuh"

namespace mrt {

#define MRT_STACK8 32          /* node groups; host checks the tree depth against it (api.cu) */

// Visiting order of child slots per ray octant: lut[c][x] = x with bit i moved to bit i ^ c.  2 KB of shared memory,
// filled once per block by the kernel (every thread of the block must call this) and handed to the traversal.
// Behind it, 256 words: slot mask -> the same bits tripled (mrt_spread3), which ANDed with Bvh8Node::leafTris gives the
// triangles of the hit leaf children without a loop over the slots.  Declare the table 4-byte aligned.
#define MRT_ORDER_LUT_BYTES (8 * 256 + 256 * 4)
__device__ __forceinline__ void bvh8_fill_order_lut(uint8_t* lut) {
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) lut[i] = (uint8_t)mrt_perm8(i & 0xffu, i >> 8);
    uint32_t* spread = reinterpret_cast<uint32_t*>(lut + 8 * 256);
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x) spread[i] = mrt_spread3(i);
    __syncthreads();
}

template <class IO, bool COUNT = false>
__device__ __forceinline__ void trace_persistent8(const float4* __restrict__ nodes8, const float4* __restrict__ tris,
                                                  uint32_t total, unsigned int* head, uint32_t refillBelow, uint32_t leafMin,
                                                  uint32_t magic, const uint8_t* orderLut, IO& io,
                                                  unsigned long long* visits = nullptr) {
    // COUNT: diagnostic build that adds up node tests and triangle tests into visits[0], visits[1] (MRT_TRACE_COUNT=1)
    uint32_t nNodes = 0, nTris = 0;
    // magic = 0x47000000 arrives as a kernel PARAMETER: were it a literal, ptxas would make it PRMT's immediate and
    // spend an extra move per plane on the byte selector instead (bvh8.h: mrt_plane)
    // orderLut: the block's shared-memory table filled by bvh8_fill_order_lut
    const uint32_t lane = threadIdx.x & 31u;
    const float tiny = 1e-30f;
    // leafMin carries the vote tuning: bits 0-7 = lanes with triangles that force a triangle step, bits 8-15 = ratio r
    // (triangle step as soon as 16 * lanes-with-triangles >= r * lanes-with-node-work; 0 = off).  A triangle step
    // costs ~0.4 node steps, so r = 6..7 picks whichever step advances more lanes per instruction (pass M: +2 %; an L1
    // prefetch of the leaf's first triangle at the end of the node step was tried there too and lost 1-20 %).
    const int triRatio = int((leafMin >> 8) & 0xffu);
    leafMin &= 0xffu;
    bool alive = false, exhausted = (total == 0);
    uint32_t tag = 0, target = MRT_NO_HIT;
    f3 o = mk3(0.f, 0.f, 0.f), d = o;
    float tmin = 0.f, tmax = 0.f;
    Ray8 r; r.ox = r.oy = r.oz = 0.f; r.invx = r.invy = r.invz = 0.f; r.negx = r.negy = r.negz = 0u; r.octInv = 0u; r.magic = magic;
    HitRec best; best.t = 0.f; best.tri = MRT_NO_HIT; best.u = best.v = 0.f;
    uint32_t gx = 0, gy = 0, tx = 0, ty = 0, tv = 0;       // tv: leafTris of the node T came from (bit -> record offset)
    const uint32_t* spreadLut = reinterpret_cast<const uint32_t*>(orderLut + 8 * 256);
    uint2 stack[MRT_STACK8];
    int sp = 0;

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
                        r.ox = o.x; r.oy = o.y; r.oz = o.z;
                        r.invx = __fdiv_rn(1.0f, fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x);
                        r.invy = __fdiv_rn(1.0f, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y);
                        r.invz = __fdiv_rn(1.0f, fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
                        r.negx = __float_as_uint(d.x) >> 31; r.negy = __float_as_uint(d.y) >> 31; r.negz = __float_as_uint(d.z) >> 31;
                        r.octInv = 7u - (r.negx | (r.negy << 1) | (r.negz << 2));
                        best.t = tmax; best.tri = target; best.u = 0.0f; best.v = 0.0f;
                        gx = 0; gy = 0x01000000u; tx = 0; ty = 0; sp = 0;       // group holding just the root (as slot 0)
                        alive = true;
                    }
                }
            }
            exhausted = base + n >= total;
        }
        if (am == 0) break;                               // queue empty and every lane idle

        for (;;) {
            const bool wantTri = alive && ty != 0;
            const bool wantNode = alive && ty == 0;        // invariant: an alive lane with T empty has inner hits in G
            const unsigned mN = __ballot_sync(0xffffffffu, wantNode);
            const unsigned mT = __ballot_sync(0xffffffffu, wantTri);
            // mN | mT = the lanes still alive: back to the refill when the warp runs thin (or has nothing left)
            const int nT = __popc(mT), nN = __popc(mN);
            if (nN + nT == 0 || (!exhausted && nN + nT < (int)refillBelow)) break;
            if (mN != 0 && nT < (int)leafMin && (triRatio == 0 || 16 * nT < triRatio * nN)) {
                if (wantNode) {
                    // nearest unvisited inner child in octant order (G keeps its hits in slot space)
                    const uint32_t pr = orderLut[(r.octInv << 8) | (gy >> 24)];
                    const uint32_t slot = uint32_t(31 - __clz(pr)) ^ r.octInv;
                    gy &= ~(0x01000000u << slot);
                    if (gy & 0xff000000u) stack[sp++] = make_uint2(gx, gy);
                    const uint32_t rel = __popc(gy & 0xffu & ((1u << slot) - 1u));
                    const float4* np = nodes8 + 5 * (size_t)(gx + rel);
                    uint32_t w[20];
                    {
                        float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3), q4 = __ldg(np + 4);
                        w[0] = __float_as_uint(q0.x); w[1] = __float_as_uint(q0.y); w[2] = __float_as_uint(q0.z); w[3] = __float_as_uint(q0.w);
                        w[4] = __float_as_uint(q1.x); w[5] = __float_as_uint(q1.y); w[6] = __float_as_uint(q1.z); w[7] = __float_as_uint(q1.w);
                        w[8] = __float_as_uint(q2.x); w[9] = __float_as_uint(q2.y); w[10] = __float_as_uint(q2.z); w[11] = __float_as_uint(q2.w);
                        w[12] = __float_as_uint(q3.x); w[13] = __float_as_uint(q3.y); w[14] = __float_as_uint(q3.z); w[15] = __float_as_uint(q3.w);
                        w[16] = __float_as_uint(q4.x); w[17] = __float_as_uint(q4.y); w[18] = __float_as_uint(q4.z); w[19] = __float_as_uint(q4.w);
                    }
                    if (COUNT) ++nNodes;
                    const uint32_t hs = bvh8_hit_slots(w, r, tmin, best.t);
                    const uint32_t imask = w[3] >> 24;
                    gx = w[4]; gy = ((hs & imask) << 24) | imask;
                    tx = w[5]; tv = w[6]; ty = spreadLut[hs & ~imask] & tv;
                    if (ty == 0 && !(gy & 0xff000000u)) {                       // nothing hit below: next group or done
                        if (sp > 0) { uint2 e = stack[--sp]; gx = e.x; gy = e.y; }
                        else { if (best.tri == MRT_NO_HIT) best.t = -1.0f; io.store(tag, target, best); alive = false; }
                    }
                }
                continue;
            }
            if (wantTri) {                                  // one triangle per step: lanes with longer lists re-vote
                const int k = __ffs(ty) - 1;
                ty &= ty - 1u;
                if (COUNT) ++nTris;
                triangle_test(tris, tx + bvh8_triangle_record(tv, uint32_t(k)), o, d, tmin, tmax, best);
                if (target != MRT_NO_HIT && best.tri != target) {                // any-hit job: occluded, done
                    ty = 0; io.store(tag, target, best); alive = false;
                } else if (ty == 0 && !(gy & 0xff000000u)) {
                    if (sp > 0) { uint2 e = stack[--sp]; gx = e.x; gy = e.y; }
                    else { if (best.tri == MRT_NO_HIT) best.t = -1.0f; io.store(tag, target, best); alive = false; }
                }
            }
        }
    }
    if (COUNT && visits) {
        nNodes = __reduce_add_sync(0xffffffffu, nNodes); nTris = __reduce_add_sync(0xffffffffu, nTris);
        if (lane == 0) { atomicAdd(visits, (unsigned long long)nNodes); atomicAdd(visits + 1, (unsigned long long)nTris); }
    }
}

}  // namespace mrt
