// bvh8.h — 8-wide BVH with 8-bit quantised child boxes: 80-byte nodes (5 x 128-bit loads), built on the host by
// collapsing the binned-SAH BVH2 of bvh.cpp, traversed by device_trace8.cuh.
//
// Replaces MPSTriangleAccelerationStructure / MPSRayIntersector (reference: renderer/Renderer.mm:456-469, 519-523,
// 545-553; closed source).  Why 8-wide: on B200 the traversal is bound by the latency of dependent node fetches
// (ncu, BVH2: long-scoreboard 16 of 23 stall cycles per issue, DRAM < 5 %); an 8-ary node cuts the dependent chain
// ~3x and packs 8 child boxes in 80 B instead of 256 B, so far more of the tree stays in L1/L2.
// Layout after Ylitie, Karras, Laine, "Efficient incoherent ray traversal on GPUs through compressed wide BVHs"
// (HPG 2017): the node stores an origin p, a power-of-two scale per axis and, per child, 6 bytes of box planes
// relative to p; children that are inner nodes are stored contiguously from childBase (slot order), triangles of
// leaf children contiguously from triBase.  The boxes are CONSERVATIVE (quantised outwards around boxes that already
// carry the builder's padding), so the result of a traversal is still min over all triangles of (t, index).
#pragma once
#include <cstdint>
#include <vector>
#include "bvh.h"

#if defined(__CUDACC__)
#define MRT_HD __host__ __device__ __forceinline__
#else
#include <cmath>
#define MRT_HD inline
#endif

namespace mrt {

struct Bvh8Node {
    float p[3];             // origin of the quantisation grid (= lower corner of the node box)
    uint8_t e[3];           // biased exponents: grid step on axis a = 2^(e[a]-127)
    uint8_t imask;          // bit s set: child slot s is an inner node
    uint32_t childBase;     // node index of the first inner child
    uint32_t triBase;       // triangle record index of the first triangle of this node's leaf children
    uint32_t leafTris;      // bit 3s+j: the leaf child in slot s holds more than j triangles (<= 3 per leaf).  The triangles of
                            // a node's leaf children lie contiguously from triBase in slot order, so triangle (s, j) is
                            // record triBase + popcount(leafTris & ((1 << (3s+j)) - 1)): no per-slot offsets, no loop
    uint32_t reserved;
    uint8_t qlo[3][8];      // lower planes x,y,z per child slot
    uint8_t qhi[3][8];      // upper planes
};
static_assert(sizeof(Bvh8Node) == 80, "wide node is 80 bytes");

struct Bvh8Result {
    std::vector<Bvh8Node> nodes;    // nodes[0] = root
    std::vector<BvhTri> tris;       // 48-byte records (bvh.h), leaf order of THIS tree
    uint32_t depth = 0;             // levels of wide nodes
    float sah_cost = 0.0f;
    double build_ms = 0.0;
    int bins = 16;                  // SAH bins per axis the binary builder used (build_bvh8 picks among candidates)
    float sample_cost = 0.0f;       // node tests + 1.2 x triangle tests per ray of the selection sample (0 = no selection)
};

// Collapse + quantise.  Same inputs as build_bvh; deterministic for any thread count.
void build_bvh8(const mrt_vertex* vertices, const uint32_t* indices, uint32_t triangle_count, uint32_t threads, Bvh8Result& out);

// Host-only check for the CPU test-suite: traverses `count` rays (8 floats each: origin, tmin, direction, tmax) through
// the wide tree with the same node decode and state machine as the CUDA kernel and compares (t, triangle) with brute
// force over all triangles.  Returns the number of rays that differ.
uint32_t bvh8_selftest(const Bvh8Result& bvh, const float* rays, uint32_t count, uint64_t* nodesVisited, uint64_t* trisTested);

// ---- node decode shared by the CUDA traversal and the host self-test --------------------------------------------
// Ray constants for the quantised slab test.  inv* = 1/d with |d| clamped away from zero (box test only).
struct Ray8 {
    float ox, oy, oz;
    float invx, invy, invz;
    uint32_t negx, negy, negz;      // direction sign per axis (1 = negative)
    uint32_t octInv;                // 7 - octant: traversal priority of slot s is s ^ octInv (highest first)
    uint32_t magic;                 // 0x47000000, kept in a REGISTER so that PRMT takes its selector as an immediate
};

MRT_HD float mrt_bits_float(uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(b);
#else
    union { uint32_t u; float f; } c; c.u = b; return c.f;
#endif
}
MRT_HD float mrt_fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return std::fmaf(a, b, c);
#endif
}
// min/max of finite values (FMNMX / FMNMX3 on sm_100a)
MRT_HD float mrt_min3(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return fminf(fminf(a, b), c);
#else
    float m = a < b ? a : b; return m < c ? m : c;
#endif
}
MRT_HD float mrt_max3(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return fmaxf(fmaxf(a, b), c);
#else
    float m = a > b ? a : b; return m > c ? m : c;
#endif
}
MRT_HD float mrt_abs(float a) {           // |a| as an operand modifier on the device (a compare + select otherwise)
#if defined(__CUDA_ARCH__)
    return fabsf(a);
#else
    return a < 0.0f ? -a : a;
#endif
}
MRT_HD float mrt_min2(float a, float b) {
#if defined(__CUDA_ARCH__)
    return fminf(a, b);
#else
    return a < b ? a : b;
#endif
}
MRT_HD float mrt_max2(float a, float b) {
#if defined(__CUDA_ARCH__)
    return fmaxf(a, b);
#else
    return a > b ? a : b;
#endif
}
// byte j of w as the float 32768 + byte: one PRMT builds the bit pattern 0x47 00 bb 00 (an int->float conversion
// would go through the quarter-rate XU pipe: ncu showed it 65 % busy with 48 I2F.U8 per node)
MRT_HD float mrt_plane(uint32_t w, int j, uint32_t magic) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(w, magic, 0x7404u | (uint32_t(j) << 4)));
#else
    return mrt_bits_float(magic | (((w >> (8 * j)) & 0xffu) << 8));
#endif
}
// byte s (0..7) of the 8-byte pair (lo, hi)
MRT_HD uint32_t mrt_byte8(uint32_t lo, uint32_t hi, uint32_t s) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(lo, hi, 0x4440u | s) & 0xffu;
#else
    return ((s < 4 ? lo : hi) >> (8 * (s & 3u))) & 0xffu;
#endif
}
// x with bit i moved to bit i ^ c (c in 0..7): the order in which a ray of octant 7 - c wants to visit child slots
MRT_HD uint32_t mrt_perm8(uint32_t x, uint32_t c) {
    if (c & 1u) x = ((x & 0xaau) >> 1) | ((x & 0x55u) << 1);
    if (c & 2u) x = ((x & 0xccu) >> 2) | ((x & 0x33u) << 2);
    if (c & 4u) x = ((x & 0xf0u) >> 4) | ((x & 0x0fu) << 4);
    return x;
}

// Children of one node hit by one ray, as a mask over child SLOTS (bit s = slot s).  w[] = the node as 20 32-bit
// words.  Non-strict comparison: boxes touching [tmin, tmax] count as hit.  Empty slots carry an inverted box.
//
// Plane k of a child lies at p + q*step; along the ray that is t = q*(step/d) + (p - o)/d.  With the byte decoded as
// F = 32768 + q the same value is F*a + (b - 32768*a), a = step/d, b = (p - o)/d: one FFMA per plane.  The constant
// c = b - 32768*a carries a rounding error of up to 2^-9 |a| (twice, once more when the guard is added), so entry
// planes use c - 2^-7 |a| and exit planes c + 2^-7 |a|: the test stays conservative at a cost of < 1 % of a grid step.
MRT_HD uint32_t bvh8_hit_slots(const uint32_t* w, const Ray8& r, float tmin, float tmax) {
    const uint32_t ew = w[3];
    const float sx = mrt_bits_float((ew & 0xffu) << 23), sy = mrt_bits_float(((ew >> 8) & 0xffu) << 23),
                sz = mrt_bits_float(((ew >> 16) & 0xffu) << 23);
    const float ax = sx * r.invx, ay = sy * r.invy, az = sz * r.invz;
    const float bx = (mrt_bits_float(w[0]) - r.ox) * r.invx, by = (mrt_bits_float(w[1]) - r.oy) * r.invy,
                bz = (mrt_bits_float(w[2]) - r.oz) * r.invz;
    const float cx = mrt_fma(-32768.0f, ax, bx), cy = mrt_fma(-32768.0f, ay, by), cz = mrt_fma(-32768.0f, az, bz);
    const float gx = mrt_abs(ax) * 0.0078125f, gy = mrt_abs(ay) * 0.0078125f, gz = mrt_abs(az) * 0.0078125f;
    const float nxc = cx - gx, fxc = cx + gx, nyc = cy - gy, fyc = cy + gy, nzc = cz - gz, fzc = cz + gz;
    const uint32_t k = r.magic;
#if defined(__CUDA_ARCH__)
    // Device form of the loop below: the miss bit of a child is the sign of tf - tn, gathered with one funnel shift per
    // child (children 7..0, so that child 0 ends in bit 0) instead of compare + select + add — 11 ALU-pipe instructions
    // fewer per node (pass X: +2 %).  Same predicate as `tn <= tf`: equal values give +0; inf - inf gives the canonical
    // positive NaN (hit, as inf <= inf); the one pair that would differ, tf = -0 with tn = +0, cannot occur — an exit
    // value is F*a + (c + |a|*2^-7) with F > 0, which is never -0, and tmax is taken as tmax + 0.
    tmax = tmax + 0.0f;
    uint32_t miss = 0;
#pragma unroll
    for (int h = 1; h >= 0; --h) {
        const uint32_t nx = r.negx ? w[14 + h] : w[8 + h], fx = r.negx ? w[8 + h] : w[14 + h];
        const uint32_t ny = r.negy ? w[16 + h] : w[10 + h], fy = r.negy ? w[10 + h] : w[16 + h];
        const uint32_t nz = r.negz ? w[18 + h] : w[12 + h], fz = r.negz ? w[12 + h] : w[18 + h];
#pragma unroll
        for (int j = 3; j >= 0; --j) {
            const float t0x = mrt_fma(mrt_plane(nx, j, k), ax, nxc), t1x = mrt_fma(mrt_plane(fx, j, k), ax, fxc);
            const float t0y = mrt_fma(mrt_plane(ny, j, k), ay, nyc), t1y = mrt_fma(mrt_plane(fy, j, k), ay, fyc);
            const float t0z = mrt_fma(mrt_plane(nz, j, k), az, nzc), t1z = mrt_fma(mrt_plane(fz, j, k), az, fzc);
            const float tn = mrt_max2(mrt_max3(t0x, t0y, t0z), tmin), tf = mrt_min2(mrt_min3(t1x, t1y, t1z), tmax);
            miss = __funnelshift_l(__float_as_uint(tf - tn), miss, 1);
        }
    }
    return ~miss & 0xffu;
#else
    uint32_t hit = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int h = 0; h < 2; ++h) {
        // words: qlo x 8+h, y 10+h, z 12+h; qhi x 14+h, y 16+h, z 18+h
        const uint32_t nx = r.negx ? w[14 + h] : w[8 + h], fx = r.negx ? w[8 + h] : w[14 + h];
        const uint32_t ny = r.negy ? w[16 + h] : w[10 + h], fy = r.negy ? w[10 + h] : w[16 + h];
        const uint32_t nz = r.negz ? w[18 + h] : w[12 + h], fz = r.negz ? w[12 + h] : w[18 + h];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int j = 0; j < 4; ++j) {
            const float t0x = mrt_fma(mrt_plane(nx, j, k), ax, nxc), t1x = mrt_fma(mrt_plane(fx, j, k), ax, fxc);
            const float t0y = mrt_fma(mrt_plane(ny, j, k), ay, nyc), t1y = mrt_fma(mrt_plane(fy, j, k), ay, fyc);
            const float t0z = mrt_fma(mrt_plane(nz, j, k), az, nzc), t1z = mrt_fma(mrt_plane(fz, j, k), az, fzc);
            // [tn, tf] clipped to [tmin, tmax] (tmin <= tmax on every live ray): one compare per child instead of three
            const float tn = mrt_max2(mrt_max3(t0x, t0y, t0z), tmin), tf = mrt_min2(mrt_min3(t1x, t1y, t1z), tmax);
            if (tn <= tf) hit |= 1u << (4 * h + j);
        }
    }
    return hit;
#endif
}

// x (8 bits) with every bit tripled: bit s -> bits 3s, 3s+1, 3s+2 (slot mask -> positions in Bvh8Node::leafTris).  The
// trace kernel reads this from a 1 KB shared-memory table instead (device_trace8.cuh).
MRT_HD uint32_t mrt_spread3(uint32_t x) {
    x = (x | (x << 8)) & 0x00F00Fu;
    x = (x | (x << 4)) & 0x0C30C3u;
    x = (x | (x << 2)) & 0x249249u;
    return x * 7u;
}
MRT_HD uint32_t mrt_popc(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
// Triangles (bits of Bvh8Node::leafTris) of the leaf children in `leafSlots`; triangle of bit k = record
// triBase + bvh8_triangle_record(leafTris, k).
MRT_HD uint32_t bvh8_leaf_triangles(uint32_t leafTris, uint32_t leafSlots) { return mrt_spread3(leafSlots) & leafTris; }
MRT_HD uint32_t bvh8_triangle_record(uint32_t leafTris, uint32_t k) { return mrt_popc(leafTris & ((1u << k) - 1u)); }
// The same as a mask over RECORDS (bit = offset from triBase) — the traversal variants of the experiments build and the
// host self-test walk this one.
MRT_HD uint32_t bvh8_triangle_mask(uint32_t leafTris, uint32_t leafSlots) {
    uint32_t t = bvh8_leaf_triangles(leafTris, leafSlots), m = 0;
    while (t) {
#if defined(__CUDA_ARCH__)
        const uint32_t k = uint32_t(__ffs(t)) - 1u;
#else
        const uint32_t k = uint32_t(__builtin_ctz(t));
#endif
        t &= t - 1u;
        m |= 1u << bvh8_triangle_record(leafTris, k);
    }
    return m;
}
// (unary triangle count << 5) | offset from triBase of the leaf child in slot s (the experiments' triangle queue)
MRT_HD uint32_t bvh8_leaf_meta(uint32_t leafTris, uint32_t s) {
    return (((leafTris >> (3u * s)) & 7u) << 5) | bvh8_triangle_record(leafTris, 3u * s);
}

}  // namespace mrt
