// device_trace8_experiments.cuh — traversal variants that were measured and did NOT become the production kernel
// (DESIGN.md 3.2): the warp-wide triangle queue (MRT_TRACE_TQ) and postponed triangles (MRT_TRACE_DEFER).  Compiled only
// into the -DMRT_EXPERIMENTS library (lib/libmrt_b200_exp.so); same contract and same bits as trace_persistent8.
#pragma once
#include "device_trace8.cuh"

namespace mrt {

// Variant with a WARP-WIDE TRIANGLE QUEUE (MRT_TRACE_TQ=1).  In the kernel above a triangle step runs with the 7-10 lanes
// that happen to hold triangles while the other ~20 wait.  Here a lane that finds leaf triangles does not test them
// itself: it appends (owner lane, triangle record) pairs to a queue in shared memory and goes on traversing; when the
// queue holds `triThresh` pairs (or nobody has node work left) the warp runs ONE triangle step in which lane i tests
// pair i — any lane tests any ray's triangle, reading the ray from shared memory — so triangle steps run with (nearly)
// all lanes.  The nearest hit of a ray lives in shared memory as the 64-bit key (ordered t, triangle index), updated
// with atomicMin: that IS the rule "smaller t, ties to the lower index".  The owner refreshes its culling bound from
// the key after every triangle step.  Price: a ray's bound tightens a few node steps later.  The answer is unchanged —
// the minimum over all triangles of (t, index) does not depend on when a triangle is tested.
#define MRT_TQ_CAP 128u                       /* pairs per warp (power of two), >= 2 x 32 */
struct TqWarp {
    float ray[7][32];                         // origin.xyz, direction.xyz, tmin of each lane's ray
    unsigned long long key[32];               // (ordered t << 32) | triangle index of the best hit so far
    float hit[3][32];                         // raw t, u, v of the key's holder
    uint32_t q[MRT_TQ_CAP];                   // owner lane << 27 | triangle record index
    uint32_t tail;                            // queue tail as the lanes reserve it (atomicAdd); mirrors the register qTail
    uint32_t pad[3];
};
// byte mask (0xff per set bit) of the low 4 bits of x
__device__ __forceinline__ uint32_t tq_expand4(uint32_t x) { return (((x & 0xfu) * 0x00204081u) & 0x01010101u) * 0xffu; }
__device__ __forceinline__ uint32_t tq_order(float t) {        // float -> uint32 with the same order; -0 counts as +0
    const uint32_t b = __float_as_uint(t + 0.0f);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float tq_unorder(uint32_t k) {
    return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu));
}

template <class IO, bool COUNT = false>
__device__ __forceinline__ void trace_persistent8q(const float4* __restrict__ nodes8, const float4* __restrict__ tris,
                                                   uint32_t total, unsigned int* head, uint32_t refillBelow, uint32_t triThresh,
                                                   uint32_t magic, const uint8_t* orderLut, TqWarp* W, IO& io,
                                                   unsigned long long* visits = nullptr) {
    uint32_t nNodes = 0, nTris = 0;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t ltMask = (1u << lane) - 1u;
    const float tiny = 1e-30f;
    bool alive = false, exhausted = (total == 0);
    uint32_t tag = 0, target = MRT_NO_HIT;
    float tmin = 0.f, tmax0 = 0.f, bestT = 0.f;
    Ray8 r; r.ox = r.oy = r.oz = 0.f; r.invx = r.invy = r.invz = 0.f; r.negx = r.negy = r.negz = 0u; r.octInv = 0u; r.magic = magic;
    uint32_t gx = 0, gy = 0, tx = 0, ty = 0, m0 = 0, m1 = 0;
    uint2 stack[MRT_STACK8];
    int sp = 0;
    uint32_t qHead = 0, qTail = 0;            // warp-uniform, monotonic; queue slot = position & (CAP - 1)
    if (lane == 0) W->tail = 0u;
    __syncwarp();
    uint32_t pendEnd = 0;                     // position after this lane's last queued pair

    // one triangle step: lane i tests pair qHead + i
    auto triangle_step = [&]() {
        __syncwarp();                                       // pairs and rays written by other lanes
        const uint32_t n = min(32u, qTail - qHead);
        bool valid = false;
        unsigned long long mykey = 0;
        uint32_t owner = 0;
        float t = 0.f, u = 0.f, v = 0.f;
        if (lane < n) {
            const uint32_t e = W->q[(qHead + lane) & (MRT_TQ_CAP - 1u)];
            owner = e >> 27;
            const float4* p = tris + 3 * (size_t)(e & 0x07ffffffu);
            const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
            const f3 o = mk3(W->ray[0][owner], W->ray[1][owner], W->ray[2][owner]);
            const f3 d = mk3(W->ray[3][owner], W->ray[4][owner], W->ray[5][owner]);
            const float tmn = W->ray[6][owner];
            if (COUNT) ++nTris;
            // same operations as triangle_test (device_trace.cuh); the range test against tmax is the key comparison
            const f3 V2 = mk3(a.x, a.y, a.z), e1 = mk3(b.x, b.y, b.z), e2 = mk3(c.x, c.y, c.z);
            const f3 pv = cross(d, e2);
            const float det = dot(e1, pv);
            const float inv = __fdiv_rn(1.0f, det);
            const f3 sv = o - V2;
            u = dot(sv, pv) * inv;
            if (u >= 0.0f && u <= 1.0f) {
                const f3 qv = cross(sv, e1);
                v = dot(d, qv) * inv;
                if (v >= 0.0f && u + v <= 1.0f) {
                    t = dot(e2, qv) * inv;
                    if (t >= tmn) {
                        valid = true;
                        mykey = ((unsigned long long)tq_order(t) << 32) | (unsigned long long)__float_as_uint(a.w);
                        atomicMin(&W->key[owner], mykey);
                    }
                }
            }
        }
        __syncwarp();
        if (valid && W->key[owner] == mykey) { W->hit[0][owner] = t; W->hit[1][owner] = u; W->hit[2][owner] = v; }
        qHead += n;
        __syncwarp();
        const unsigned long long kk = W->key[lane];
        bestT = tq_unorder((uint32_t)(kk >> 32));
        if (target != MRT_NO_HIT && (uint32_t)kk != target) { gy &= 0x00ffffffu; sp = 0; }   // any-hit ray occluded: stop traversing
    };

    for (;;) {
        // ---- refill idle lanes (a dead lane has no pairs in the queue) ----
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
                uint32_t it = base + __popc(need & ltMask);
                if (it < total) {
                    TraceJob job;
                    if (io.load(it, job)) {
                        tag = job.tag; tmin = job.tmin; tmax0 = job.tmax; target = job.target;
                        const f3 o = job.o, d = job.d;
                        W->ray[0][lane] = o.x; W->ray[1][lane] = o.y; W->ray[2][lane] = o.z;
                        W->ray[3][lane] = d.x; W->ray[4][lane] = d.y; W->ray[5][lane] = d.z; W->ray[6][lane] = tmin;
                        W->key[lane] = ((unsigned long long)tq_order(tmax0) << 32) | (unsigned long long)target;
                        r.ox = o.x; r.oy = o.y; r.oz = o.z;
                        r.invx = __fdiv_rn(1.0f, fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x);
                        r.invy = __fdiv_rn(1.0f, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y);
                        r.invz = __fdiv_rn(1.0f, fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
                        r.negx = __float_as_uint(d.x) >> 31; r.negy = __float_as_uint(d.y) >> 31; r.negz = __float_as_uint(d.z) >> 31;
                        r.octInv = 7u - (r.negx | (r.negy << 1) | (r.negz << 2));
                        bestT = tq_unorder(tq_order(tmax0));
                        gx = 0; gy = 0x01000000u; tx = 0; ty = 0; sp = 0;       // group holding just the root (as slot 0)
                        pendEnd = qHead;
                        alive = true;
                    }
                }
            }
            exhausted = base + n >= total;
        }
        if (am == 0) break;                               // queue of rays empty, every lane idle (and no pairs pending)
        __syncwarp();
        unsigned aliveMask = __ballot_sync(0xffffffffu, alive);

        for (;;) {
            // ---- choose: a triangle step when enough pairs wait or nobody can take a node step ----
            const bool wantNode = alive && (gy & 0xff000000u) != 0u;
            const unsigned mN = __ballot_sync(0xffffffffu, wantNode);
            const uint32_t qsz = qTail - qHead;
            const unsigned mW = aliveMask & ~mN;                                          // lanes waiting for their last pairs
            if (qsz >= (triThresh & 0xffu) || (qsz != 0 && (mN == 0 || (uint32_t)__popc(mW) >= (triThresh >> 8)))) {
                triangle_step();
            } else if (mN != 0) {
                if (wantNode) {
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
                    const uint32_t hs = bvh8_hit_slots(w, r, tmin, bestT);
                    const uint32_t imask = w[3] >> 24;
                    gx = w[4]; gy = ((hs & imask) << 24) | imask;
                    tx = w[5]; ty = hs & ~imask;                               // ty: hit LEAF SLOTS of this node (not yet triangles)
                    m0 = w[6]; m1 = 0u;                                        // m0: leafTris of the node
                    if (!(gy & 0xff000000u) && sp > 0) { uint2 e = stack[--sp]; gx = e.x; gy = e.y; }   // G always holds the next inner work
                }
                // ---- queue the triangles just found: count them without a loop (a leaf's meta byte carries its triangle
                // count in unary in bits 5-7), reserve queue positions with one shared-memory atomic per lane, then one
                // pass over the hit leaf slots writes the pairs ----
                {
                    const uint32_t c = (uint32_t)__popc(bvh8_leaf_triangles(m0, ty));
                    const uint32_t totalNew = __reduce_add_sync(0xffffffffu, c);
                    if (qTail - qHead + totalNew <= MRT_TQ_CAP) {
                        if (c != 0u) {
                            uint32_t pos = atomicAdd(&W->tail, c);
                            pendEnd = pos + c;
                            const uint32_t laneTag = lane << 27;
                            do {
                                const uint32_t sl = uint32_t(__ffs(ty)) - 1u;
                                ty &= ty - 1u;
                                const uint32_t mb = bvh8_leaf_meta(m0, sl);
                                const uint32_t e = laneTag | (tx + (mb & 0x1fu));
                                W->q[pos & (MRT_TQ_CAP - 1u)] = e;
                                if (mb & 0x40u) W->q[(pos + 1u) & (MRT_TQ_CAP - 1u)] = e + 1u;
                                if (mb & 0x80u) W->q[(pos + 2u) & (MRT_TQ_CAP - 1u)] = e + 2u;
                                pos += __popc(mb >> 5);
                            } while (ty != 0u);
                        }
                        qTail += totalNew;
                    } else {
                        ty = bvh8_triangle_mask(m0, ty);
                        for (;;) {                          // one pair per lane per round, draining the queue when it is full
                            const unsigned mP = __ballot_sync(0xffffffffu, ty != 0u);
                            if (mP == 0) break;
                            if (qTail - qHead > MRT_TQ_CAP - 32u) { triangle_step(); continue; }
                            if (ty != 0u) {
                                const uint32_t k = uint32_t(__ffs(ty)) - 1u;
                                ty &= ty - 1u;
                                const uint32_t pos = qTail + __popc(mP & ltMask);
                                W->q[pos & (MRT_TQ_CAP - 1u)] = (lane << 27) | (tx + k);
                                pendEnd = pos + 1u;
                            }
                            qTail += __popc(mP);
                        }
                        if (lane == 0) W->tail = qTail;
                    }
                }
                __syncwarp();
            }
            // ---- retire: a ray ends when it has no node work and no pair waiting ----
            if (alive && !(gy & 0xff000000u) && (int32_t)(pendEnd - qHead) <= 0) {
                const uint32_t bestTri = (uint32_t)W->key[lane];
                HitRec best; best.tri = bestTri;
                if (bestTri == target) { best.t = (target == MRT_NO_HIT) ? -1.0f : tmax0; best.u = 0.0f; best.v = 0.0f; }
                else { best.t = W->hit[0][lane]; best.u = W->hit[1][lane]; best.v = W->hit[2][lane]; }
                io.store(tag, target, best);
                alive = false;
            }
            aliveMask = __ballot_sync(0xffffffffu, alive);
            if (aliveMask == 0 || (!exhausted && __popc(aliveMask) < (int)refillBelow)) break;
        }
    }
    if (COUNT && visits) {
        nNodes = __reduce_add_sync(0xffffffffu, nNodes); nTris = __reduce_add_sync(0xffffffffu, nTris);
        if (lane == 0) { atomicAdd(visits, (unsigned long long)nNodes); atomicAdd(visits + 1, (unsigned long long)nTris); }
    }
}

// Variant with POSTPONED triangles (MRT_TRACE_DEFER=1): a lane that finds leaf triangles keeps descending while it has
// inner work, parking up to MRT_TQUEUE triangle groups; the warp takes a triangle step when `triThresh` lanes hold
// triangles (or `leafMin` lanes have nothing else left).  Triangle steps then run with more lanes and lanes holding
// triangles no longer idle through node steps; the price is that the nearest-hit bound tightens later, so a ray may
// test a few nodes the ordered traversal would have culled.  The answer is unchanged: it is the minimum over all
// triangles of (t, index), whatever the order.
#define MRT_TQUEUE 4u

template <class IO, bool COUNT = false>
__device__ __forceinline__ void trace_persistent8d(const float4* __restrict__ nodes8, const float4* __restrict__ tris,
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
    const int triThresh = int((leafMin >> 18) & 0x3fu) ? int((leafMin >> 18) & 0x3fu) : 16;    // bits 18-23: see below
    leafMin &= 0xffu;
    bool alive = false, exhausted = (total == 0);
    uint32_t tag = 0, target = MRT_NO_HIT;
    f3 o = mk3(0.f, 0.f, 0.f), d = o;
    float tmin = 0.f, tmax = 0.f;
    Ray8 r; r.ox = r.oy = r.oz = 0.f; r.invx = r.invy = r.invz = 0.f; r.negx = r.negy = r.negz = 0u; r.octInv = 0u; r.magic = magic;
    HitRec best; best.t = 0.f; best.tri = MRT_NO_HIT; best.u = best.v = 0.f;
    uint32_t gx = 0, gy = 0, tx = 0, ty = 0;
    uint2 stack[MRT_STACK8];
    int sp = 0;
    uint2 tq[MRT_TQUEUE];                                  // postponed triangle groups, oldest (nearest) first
    uint32_t th = 0, tn = 0;                               // queue head index, entries queued

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
                        gx = 0; gy = 0x01000000u; tx = 0; ty = 0; sp = 0; th = 0; tn = 0;       // group holding just the root (as slot 0)
                        alive = true;
                    }
                }
            }
            exhausted = base + n >= total;
        }
        if (am == 0) break;                               // queue empty and every lane idle

        for (;;) {
            // a lane holding triangles keeps taking node steps while it has inner work and room to postpone another
            // group; lanes with nothing but triangles left are the ones a triangle step must not keep waiting
            const bool wantTri = alive && ty != 0;
            const bool wantNode = alive && (gy & 0xff000000u) != 0u && tn < MRT_TQUEUE;
            const unsigned mN = __ballot_sync(0xffffffffu, wantNode);
            const unsigned mT = __ballot_sync(0xffffffffu, wantTri);
            const unsigned mO = __ballot_sync(0xffffffffu, wantTri && !wantNode);
            if ((mN | mT) == 0) break;
            const int nT = __popc(mT), nN = __popc(mN), nO = __popc(mO);
            if (mN != 0 && nO < (int)leafMin && nT < triThresh && (triRatio == 0 || 16 * nO < triRatio * nN)) {
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
                    if (!(gy & 0xff000000u) && sp > 0) { uint2 e = stack[--sp]; gx = e.x; gy = e.y; }   // G always holds the next inner work
                    const uint32_t nt = bvh8_triangle_mask(w[6], hs & ~imask);
                    if (nt != 0u) {
                        if (ty == 0u) { tx = w[5]; ty = nt; }
                        else { tq[(th + tn) % MRT_TQUEUE] = make_uint2(w[5], nt); ++tn; }
                    }
                    if (ty == 0u && !(gy & 0xff000000u)) {                      // nothing left anywhere: done
                        if (best.tri == MRT_NO_HIT) best.t = -1.0f;
                        io.store(tag, target, best); alive = false;
                    }
                }
                continue;
            }
            if (wantTri) {                                  // one triangle per step: lanes with longer lists re-vote
                const int k = __ffs(ty) - 1;
                ty &= ty - 1u;
                if (COUNT) ++nTris;
                triangle_test(tris, tx + uint32_t(k), o, d, tmin, tmax, best);
                if (target != MRT_NO_HIT && best.tri != target) {                // any-hit job: occluded, done
                    ty = 0; io.store(tag, target, best); alive = false;
                } else if (ty == 0u) {
                    if (tn != 0u) { uint2 e = tq[th]; tx = e.x; ty = e.y; th = (th + 1u) % MRT_TQUEUE; --tn; }
                    else if (!(gy & 0xff000000u)) {
                        if (best.tri == MRT_NO_HIT) best.t = -1.0f;
                        io.store(tag, target, best); alive = false;
                    }
                }
            }
            const unsigned a2 = __ballot_sync(0xffffffffu, alive);
            if (a2 == 0 || (!exhausted && __popc(a2) < (int)refillBelow)) break;
        }
    }
    if (COUNT && visits) {
        nNodes = __reduce_add_sync(0xffffffffu, nNodes); nTris = __reduce_add_sync(0xffffffffu, nTris);
        if (lane == 0) { atomicAdd(visits, (unsigned long long)nNodes); atomicAdd(visits + 1, (unsigned long long)nTris); }
    }
}

}  // namespace mrt
