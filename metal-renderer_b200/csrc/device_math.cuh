// device_math.cuh — float32 arithmetic of the path-tracing kernels (sm_100a).
//
// The reference's kernels (renderer/Shaders.metal, KernelHelpers.h, Raytracing.h:168-223) use MSL built-ins
// normalize / dot / cross / length / reflect / clamp / mix / sqrt / sin / cos under Metal fast-math, i.e. with
// no defined bits.  This file fixes ONE evaluation of each — IEEE binary32, round-to-nearest, fused multiply-add
// only where __fmaf_rn is written — and the whole translation unit is compiled with -fmad=false so nvcc never
// contracts anything else.  The CPU checker restates the same sequence, which is what makes hit IDs and
// radiance comparable bit for bit between CPU and GPU.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mrt {

struct f2 { float x, y; };
struct f3 { float x, y, z; };

__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }

// a*s + b with one rounding per component
__device__ __forceinline__ f3 madd(f3 a, float s, f3 b) {
    return mk3(__fmaf_rn(a.x, s, b.x), __fmaf_rn(a.y, s, b.y), __fmaf_rn(a.z, s, b.z));
}
// dot(): fma chain x, y, z
__device__ __forceinline__ float dot(f3 a, f3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x)); }
// cross(): fma(first product, -(second product rounded))
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return mk3(__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ float length(f3 a) { return __fsqrt_rn(dot(a, a)); }
// normalize(): multiply by the IEEE reciprocal of the IEEE square root (no rsqrt approximation)
__device__ __forceinline__ f3 normalize(f3 a) { float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot(a, a))); return a * inv; }
// reflect(i, n) = i - 2*dot(n,i)*n
__device__ __forceinline__ f3 reflect(f3 i, f3 n) { float k = 2.0f * dot(n, i); return madd(n, -k, i); }
// clamp(): compare-select, NaN passes through
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
// mix(a,b,t) = a + (b-a)*t
__device__ __forceinline__ float mixf(float a, float b, float t) { return a + (b - a) * t; }

// sin/cos of the cosine-hemisphere azimuth (Raytracing.h:215): Cody-Waite reduction by pi/2 in three parts and
// the classic single-precision minimax polynomials, each step an explicit fma.  ~1 ulp on [0, 2*pi).
__device__ __forceinline__ void sincos_pinned(float x, float* s_out, float* c_out) {
    const float TWO_OVER_PI = 0.636619772367581343f;
    const float PIO2_HI = 1.57079625129699707031f;
    const float PIO2_MID = 7.54978941586159635335e-08f;
    const float PIO2_LO = 5.39030285815811590799e-15f;
    float j = rintf(x * TWO_OVER_PI);
    float r = __fmaf_rn(-j, PIO2_HI, x);
    r = __fmaf_rn(-j, PIO2_MID, r);
    r = __fmaf_rn(-j, PIO2_LO, r);
    float r2 = r * r;
    float ps = __fmaf_rn(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = __fmaf_rn(r2, ps, -1.6666654611e-1f);
    float sr = __fmaf_rn(r * r2, ps, r);
    float pc = __fmaf_rn(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = __fmaf_rn(r2, pc, 4.166664568298827e-2f);
    float cr = __fmaf_rn(r2 * r2, pc, __fmaf_rn(r2, -0.5f, 1.0f));
    int q = (int)j & 3;
    float s = (q & 1) ? cr : sr;
    float c = (q & 1) ? sr : cr;
    if (q == 1 || q == 2) c = -c;
    if (q >= 2) s = -s;
    *s_out = s;
    *c_out = c;
}

}  // namespace mrt
