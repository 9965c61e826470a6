// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_abi.h).
//
// orc_math.h — the float32 arithmetic the oracle is defined in.
//
// The reference's kernels ran under Metal fast-math, so it never had reproducible bits; the oracle
// therefore PINS one evaluation of every MSL built-in the path uses (SURVEY.md §8c lists them):
// IEEE-754 binary32, round-to-nearest-even, no implicit contraction (build with -ffp-contract=off),
// fused multiply-add only where fmaf() is written.  The CUDA product restates the same sequence with
// the same explicit fmaf()s (nvcc -fmad=false), which is what makes hit IDs and radiance bit-comparable.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

struct f2 { float x, y; };
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };

static inline f3 ld3(const float* p) { return f3{p[0], p[1], p[2]}; }
static inline void st3(float* p, f3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }

static inline f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
static inline f3 operator-(f3 a) { return f3{-a.x, -a.y, -a.z}; }

// a*s + b, one rounding per component
static inline f3 madd(f3 a, float s, f3 b) {
    return f3{std::fmaf(a.x, s, b.x), std::fmaf(a.y, s, b.y), std::fmaf(a.z, s, b.z)};
}
// MSL dot(): fma chain x, then y, then z
static inline float dot(f3 a, f3 b) { return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x)); }
// MSL cross(): each component = fma(first product, -(second product rounded))
static inline f3 cross(f3 a, f3 b) {
    return f3{std::fmaf(a.y, b.z, -(a.z * b.y)),
              std::fmaf(a.z, b.x, -(a.x * b.z)),
              std::fmaf(a.x, b.y, -(a.y * b.x))};
}
static inline float length(f3 a) { return std::sqrt(dot(a, a)); }
// MSL normalize(): multiply by the IEEE reciprocal of the IEEE square root (no rsqrt)
static inline f3 normalize(f3 a) { float inv = 1.0f / std::sqrt(dot(a, a)); return a * inv; }
// MSL reflect(i, n) = i - 2*dot(n,i)*n
static inline f3 reflect(f3 i, f3 n) { float k = 2.0f * dot(n, i); return madd(n, -k, i); }
// MSL clamp(): compare-select, NaN passes through
static inline float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
// MSL mix(a,b,t) = a + (b-a)*t
static inline float mixf(float a, float b, float t) { return a + (b - a) * t; }

// sin/cos for the cosine-hemisphere azimuth (Raytracing.h:215).  MSL's precision there is
// implementation-defined; the oracle pins a Cody-Waite reduction by pi/2 and the classic single-precision
// minimax polynomials (Cephes sinf/cosf coefficients), every step an explicit fmaf.  Accurate to ~1 ulp
// on [0, 2*pi) (tests compare with libm).
static inline void sincos_pinned(float x, float* s_out, float* c_out) {
    const float TWO_OVER_PI = 0.636619772367581343f;
    const float PIO2_HI = 1.57079625129699707031f;       // pi/2 split in three
    const float PIO2_MID = 7.54978941586159635335e-08f;
    const float PIO2_LO = 5.39030285815811590799e-15f;
    float j = std::nearbyintf(x * TWO_OVER_PI);
    float r = std::fmaf(-j, PIO2_HI, x);
    r = std::fmaf(-j, PIO2_MID, r);
    r = std::fmaf(-j, PIO2_LO, r);
    float r2 = r * r;
    // sin(r) ~ r + r^3*(S1 + r2*(S2 + r2*S3))
    float ps = std::fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = std::fmaf(r2, ps, -1.6666654611e-1f);
    float sr = std::fmaf(r * r2, ps, r);
    // cos(r) ~ 1 - r2/2 + r2^2*(C1 + r2*(C2 + r2*C3))
    float pc = std::fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = std::fmaf(r2, pc, 4.166664568298827e-2f);
    float cr = std::fmaf(r2 * r2, pc, std::fmaf(r2, -0.5f, 1.0f));
    int q = (int)j & 3;
    float s = (q & 1) ? cr : sr;
    float c = (q & 1) ? sr : cr;
    if (q == 1 || q == 2) c = -c;
    if (q >= 2) s = -s;
    *s_out = s;
    *c_out = c;
}

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

}  // namespace orc
