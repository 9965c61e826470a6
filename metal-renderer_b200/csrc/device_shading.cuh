// device_shading.cuh — sm_100a device code for the shading half of the per-pixel pipeline:
// rayGenerator, intersectionHandler, lightSamplingHandler, accumulateImage
// (reference: renderer/Shaders.metal:75-249) and the helpers they call (renderer/KernelHelpers.h:7-190,
// renderer/Raytracing.h:168-223, renderer/Spectrum.h).  Written against register-resident path state so the
// same functions serve the fused megakernel and the per-stage kernels that keep the reference's buffers.
#pragma once
#include "device_math.cuh"
#include "mrt.h"

namespace mrt {

#define MRT_INV_PI (1.0f / MRT_PI)     /* "(1.0 / PI)", KernelHelpers.h:78 — folded in float */

// Scene as the kernels read it: 128-bit aligned records repacked from the ABI arrays at mrt_set_scene.
struct DeviceScene {
    const float4* nodes;        // binary tree: 4 per node (bvh.h)
    const float4* tris;         //              3 per triangle, leaf order of the binary tree
    const float4* nodes8;       // 8-wide tree: 5 per node (bvh8.h)
    const float4* tris8;        //              3 per triangle, leaf order of the wide tree
    const float4* shadeTris;    // 5 per triangle, original order: p0 p1 p2 n0 n1 n2 | materialIndex lightTriangleIndex
    const float4* materials;    // 2 per material: diffuse.rgb emissive.r | emissive.gb ior type
    const float4* lights;       // 6 per light triangle (count + 1 sentinel)
    const float* lightCdf;      // count + 1
    const uint32_t* lightGuide; // guide[b] = index selected at xi = b / buckets, b = 0..buckets (guide[buckets] = count); or null
    uint32_t lightGuideBuckets; // power of two, 0 = no table (few lights)
    uint32_t lightCount;
    uint32_t triangleCount;
    // optics extension (NOT in the reference, off by default — see mrt_set_extensions in mrt.h)
    const float2* lightAlias;   // per light triangle: (threshold, alias index as bits) — Walker/Vose table of the light pdfs
    uint32_t extFlags;          // MRT_EXT_* bits; 0 = the reference's behaviour, the EXT kernel variants are never launched
    float dispersion;           // MRT_EXT_DISPERSION: ior of channel c = material.ior + dispersion * (c - 1)
};

struct Vtx { f3 v, n; };
struct Mat { f3 diffuse, emissive; float ior; uint32_t type; };
struct Light { f3 emissive; Vtx v1, v2, v3; float area, pdf; uint32_t index; };
struct ShadeTri { Vtx t0, t1, t2; uint32_t material, light; };

// Path state of one pixel-sample = the live fields of the reference's Ray (Raytracing.h:54-69).
struct PathState {
    f3 origin, direction;
    f3 throughput, radiance;
    float pdf, prevDiffuse, bounce, ior;     // Ray.params x y z w
};
// Live fields of LightSamplingRay (Raytracing.h:71-83); origin is shared with the continued path.
struct ShadowRay { f3 direction; f3 throughput; uint32_t targetIndex; };

struct HitRec { float t; uint32_t tri; float u, v; };

__device__ __forceinline__ ShadeTri load_shade_tri(const DeviceScene& S, uint32_t tri) {
    const float4* p = S.shadeTris + 5 * (size_t)tri;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    ShadeTri r;
    r.t0.v = mk3(a.x, a.y, a.z); r.t1.v = mk3(a.w, b.x, b.y); r.t2.v = mk3(b.z, b.w, c.x);
    r.t0.n = mk3(c.y, c.z, c.w); r.t1.n = mk3(d.x, d.y, d.z); r.t2.n = mk3(d.w, e.x, e.y);
    r.material = __float_as_uint(e.z); r.light = __float_as_uint(e.w);
    return r;
}
__device__ __forceinline__ Mat load_material(const DeviceScene& S, uint32_t m) {
    float4 a = __ldg(S.materials + 2 * (size_t)m), b = __ldg(S.materials + 2 * (size_t)m + 1);
    Mat r; r.diffuse = mk3(a.x, a.y, a.z); r.emissive = mk3(a.w, b.x, b.y); r.ior = b.z; r.type = __float_as_uint(b.w);
    return r;
}
__device__ __forceinline__ Light load_light(const DeviceScene& S, uint32_t i) {
    const float4* p = S.lights + 6 * (size_t)i;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4), f = __ldg(p + 5);
    Light L;
    L.emissive = mk3(a.x, a.y, a.z); L.area = a.w;
    L.v1.v = mk3(b.x, b.y, b.z); L.pdf = b.w;
    L.v1.n = mk3(c.x, c.y, c.z); L.index = __float_as_uint(c.w);
    L.v2.v = mk3(d.x, d.y, d.z); L.v2.n = mk3(d.w, e.x, e.y);
    L.v3.v = mk3(e.z, e.w, f.x); L.v3.n = mk3(f.y, f.z, f.w);
    return L;
}

// ---- Raytracing.h:168-223 ------------------------------------------------------------------------------
__device__ __forceinline__ float triangleSamplePDF(float area, float cosTheta, float distanceToSample) {   // :168-171
    return __fdiv_rn(distanceToSample * distanceToSample, area * cosTheta);
}
__device__ __forceinline__ float balanceHeuristic(float fPdf, float gPdf) {                                // :173-178
    float f2_ = fPdf * fPdf, g2_ = gPdf * gPdf;
    return __fdiv_rn(f2_, f2_ + g2_);
}
__device__ __forceinline__ f3 barycentric(float sx, float sy) {                                            // :182-187
    float r1 = __fsqrt_rn(sx), r2 = sy;
    return mk3(1.0f - r1, r1 * (1.0f - r2), r1 * r2);
}
__device__ __forceinline__ void buildOrthonormalBasis(f3 n, f3& u, f3& v) {                                // :189-205
    if (n.z < 0.0f) {
        float a = __fdiv_rn(1.0f, 1.0f - n.z);
        float b = n.x * n.y * a;
        u = mk3(1.0f - n.x * n.x * a, -b, n.x);
        v = mk3(b, n.y * n.y * a - 1.0f, -n.y);
    } else {
        float a = __fdiv_rn(1.0f, 1.0f + n.z);
        float b = -n.x * n.y * a;
        u = mk3(1.0f - n.x * n.x * a, b, -n.x);
        v = mk3(b, 1.0f - n.y * n.y * a, -n.y);
    }
}
__device__ __forceinline__ f3 alignWithNormal(f3 n, float cosTheta, float phi) {                           // :207-216
    float sinTheta = __fsqrt_rn(1.0f - cosTheta * cosTheta);
    f3 u, v;
    buildOrthonormalBasis(n, u, v);
    float s, c;
    sincos_pinned(phi, &s, &c);
    return (u * c + v * s) * sinTheta + n * cosTheta;
}
__device__ __forceinline__ f3 generateDiffuseBounce(float sx, float sy, f3 n) {                            // :218-223
    float cosTheta = __fsqrt_rn(sy);
    float phi = sx * MRT_PI * 2.0f;
    return alignWithNormal(n, cosTheta, phi);
}

// ---- KernelHelpers.h -----------------------------------------------------------------------------------
__device__ __forceinline__ float fresnel(f3 n, f3 i, float etaOut, float etaIn) {                          // :7-21
    float result = 1.0f;
    float etaScale = __fdiv_rn(etaOut, etaIn);
    float cosThetaI = clampf(dot(n, i), -1.0f, 1.0f);
    float sinThetaTSquared = (etaScale * etaScale) * (1.0f - cosThetaI * cosThetaI);
    if (sinThetaTSquared < 1.0f) {
        float cosThetaT = __fsqrt_rn(1.0f - sinThetaTSquared);
        float rS = __fdiv_rn(etaIn * cosThetaI - etaOut * cosThetaT, etaIn * cosThetaI + etaOut * cosThetaT);
        float rP = __fdiv_rn(etaIn * cosThetaT - etaOut * cosThetaI, etaIn * cosThetaT + etaOut * cosThetaI);
        result = 0.5f * (rS * rS + rP * rP);
    }
    return result;
}
__device__ __forceinline__ Vtx interpolate(const Vtx& t0, const Vtx& t1, const Vtx& t2, f3 uvw) {          // :23-37
    Vtx r;
    r.v = madd(t2.v, uvw.z, madd(t1.v, uvw.y, t0.v * uvw.x));
    r.n = normalize(madd(t2.n, uvw.z, madd(t1.n, uvw.y, t0.n * uvw.x)));
    return r;
}
__device__ __forceinline__ Vtx interpolate_uv(const Vtx& t0, const Vtx& t1, const Vtx& t2, float u, float v) {  // :39-47
    return interpolate(t0, t1, t2, mk3(u, v, 1.0f - u - v));
}
// :49-54 — the reference scans the exclusive CDF linearly; a binary search over the same predicate
// (cdf[index+1] <= xi) returns the same index, including the zero-area sentinel when xi >= sum(pdf).
// With many lights (config 5: 2048) the 11 dependent loads of that search were the longest chain of the shade kernel,
// so the host also stores, for `buckets` equal slices of [0,1), the index selected at the start of each slice: the index
// for xi lies between the entries of its slice and the next (the predicate is monotone in xi), and the search starts
// from those bounds — same predicate, same answer, 0-2 steps.
__device__ __forceinline__ uint32_t selectLightTriangle(float xi, const float* cdf, uint32_t trianglesCount,
                                                        const uint32_t* guide = nullptr, uint32_t buckets = 0) {
    uint32_t lo = 0, hi = trianglesCount;
    if (buckets != 0u && xi >= 0.0f && xi < 1.0f) {
        const uint32_t b = min(uint32_t(xi * float(buckets)), buckets - 1u);
        lo = __ldg(guide + b); hi = __ldg(guide + b + 1u);
    }
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(cdf + mid + 1) <= xi) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// ---- optics extension (no counterpart in the reference; restated in oracle/orc_shading.h under the same names) -----
// Walker/Vose alias selection with ONE uniform number: slot = floor(xi * n), the fraction decides slot vs its alias.
// Selects light i with probability pdf[i] (the table is built from the same pdfs, scene.cpp build_alias_table), in two
// dependent loads whatever the number of lights; the index differs from the CDF inversion's, so images differ sample by
// sample and agree in expectation.  Never returns the sentinel.
__device__ __forceinline__ uint32_t selectLightAlias(float xi, const float2* table, uint32_t trianglesCount) {
    if (trianglesCount == 0u) return 0u;
    const float s = fminf(fmaxf(xi, 0.0f), 1.0f) * float(trianglesCount);
    const uint32_t slot = min(uint32_t(s), trianglesCount - 1u);
    const float2 e = __ldg(table + slot);
    return (s - float(slot)) < e.x ? slot : __float_as_uint(e.y);
}
// Smooth dielectric with Snell refraction.  The reference's branch (KernelHelpers.h:151-166) continues straight through
// and keeps a cosine on the reflected lobe; here: the side is taken from the sign of n.wI (n is the outward shading
// normal), the far side's index is the material's when entering and air's (the generator's 1.00029, Shaders.metal:99) when
// leaving, the near side's is the one the ray carries (Ray.params.w), total internal reflection reflects, both lobes are
// delta lobes with bsdf/pdf = 1 and the transmitted radiance is scaled by (etaI/etaT)^2 (radiance transport from the
// camera).  No stack of nested media.  Dispersion: a transmitted path entering the medium keeps ONE of its live
// throughput channels (chosen uniformly with noise.x, weight = number of live channels) and refracts with that channel's
// index; it carries that index until it leaves.  `channelScale` multiplies the throughput.  With
// MRT_EXT_CONTINUOUS_SPECTRUM the index is taken at a uniformly random place inside the chosen band (noise.w, unused at a
// dielectric otherwise) instead of at its centre: the refraction angle is sampled over a continuum of wavelengths
// although the radiometry stays 3 samples wide.
#define MRT_AIR_IOR 1.00029f
__device__ __forceinline__ f3 refractiveBounce(const Mat& material, const DeviceScene& S, f3 wI, float rayIor, f3 throughput, f3 n,
                                               float4 noiseSample, float& bsdf, float& pdf, float& ior, f3& channelScale) {
    const float cosN = dot(n, -wI);
    const bool entering = cosN > 0.0f;
    const f3 nn = entering ? n : -n;
    const float etaI = rayIor;
    float etaT = entering ? material.ior : MRT_AIR_IOR;
    const float fI = fresnel(nn, -wI, etaI, etaT);
    channelScale = mk3(1.0f, 1.0f, 1.0f);
    ior = rayIor;
    if (fI < noiseSample.y) {
        if (entering && (S.extFlags & MRT_EXT_DISPERSION) && S.dispersion != 0.0f) {
            const bool a0 = throughput.x != 0.0f, a1 = throughput.y != 0.0f, a2 = throughput.z != 0.0f;
            const uint32_t k = uint32_t(a0) + uint32_t(a1) + uint32_t(a2);
            if (k > 0u) {
                const uint32_t rank = min(uint32_t(noiseSample.x * float(k)), k - 1u);   // the rank-th live channel
                uint32_t seen = 0u, c = 3u;
                if (a0) { if (seen == rank) c = 0u; ++seen; }
                if (a1) { if (seen == rank && c == 3u) c = 1u; ++seen; }
                if (a2) { if (seen == rank && c == 3u) c = 2u; ++seen; }
                const float w = float(k);
                channelScale = mk3(c == 0u ? w : 0.0f, c == 1u ? w : 0.0f, c == 2u ? w : 0.0f);
                float band = float(c) - 1.0f;                               // -1, 0, +1: the centre of band c
                if (S.extFlags & MRT_EXT_CONTINUOUS_SPECTRUM) band += noiseSample.w - 0.5f;   // anywhere inside the band
                etaT = material.ior + S.dispersion * band;
            }
        }
        const float eta = __fdiv_rn(etaI, etaT);
        const float cosI = clampf(dot(nn, -wI), 0.0f, 1.0f);
        const float sin2T = (eta * eta) * (1.0f - cosI * cosI);
        if (sin2T < 1.0f) {
            const float cosT = __fsqrt_rn(1.0f - sin2T);
            ior = etaT; bsdf = eta * eta; pdf = 1.0f;
            return normalize(wI * eta + nn * (eta * cosI - cosT));
        }
        channelScale = mk3(1.0f, 1.0f, 1.0f);       // this channel is totally reflected: no selection was needed
    }
    bsdf = 1.0f; pdf = 1.0f;
    return reflect(wI, nn);
}

__device__ __forceinline__ f2 mirrorLobe(f3 wI, f3 wO, f3 n, float cosTheta) {                             // :67-68, 83-84, 100-101
    bool isMirrorDirection = fabsf(dot(reflect(wI, n), wO) - 1.0f) < MRT_ANGLE_EPSILON;
    f2 r; r.x = isMirrorDirection ? cosTheta : 0.0f; r.y = 1.0f;
    return r;
}
__device__ __forceinline__ f2 sampleMaterial(const Mat& material, f3 wI, f3 wO, f3 n, float4 noiseSample) {   // :56-114
    f2 result;
    float cosTheta = dot(wO, n);
    switch (material.type) {
        case MRT_MATERIAL_MIRROR:
            result = mirrorLobe(wI, wO, n, cosTheta);
            break;
        case MRT_MATERIAL_SMOOTH_PLASTIC: {
            float fI = fresnel(n, -wI, 1.0f, material.ior);
            if (fI < noiseSample.y) { result.x = MRT_INV_PI * cosTheta; result.y = result.x; }
            else result = mirrorLobe(wI, wO, n, cosTheta);
            break;
        }
        case MRT_MATERIAL_SMOOTH_DIELECTRIC: {
            float fI = fresnel(n, -wI, 1.0f, material.ior);
            if (fI < noiseSample.y) { result.x = 0.0f; result.y = 0.0f; }
            else result = mirrorLobe(wI, wO, n, cosTheta);
            break;
        }
        default:
            result.x = MRT_INV_PI * cosTheta; result.y = result.x;
            break;
    }
    return result;
}
__device__ __forceinline__ f3 generateNextBounce(const Mat& material, f3 rayDirection, float rayIor, f3 n,      // :116-179
                                                 float4 noiseSample, float& bsdf, float& pdf, float& ior) {
    f3 wI = rayDirection, wO;
    float currentIoR = rayIor;
    ior = currentIoR;
    bool diffuseLobe = true;
    switch (material.type) {
        case MRT_MATERIAL_MIRROR:
            diffuseLobe = false;
            break;
        case MRT_MATERIAL_SMOOTH_PLASTIC: {
            float fI = fresnel(n, -wI, currentIoR, material.ior);
            diffuseLobe = fI < noiseSample.y;
            break;
        }
        case MRT_MATERIAL_SMOOTH_DIELECTRIC: {
            float fI = fresnel(n, -wI, currentIoR, material.ior);
            if (fI < noiseSample.y) {           // straight through, no refraction (:156-158)
                ior = material.ior; bsdf = 1.0f; pdf = 1.0f;
                return wI;
            }
            diffuseLobe = false;
            break;
        }
        default:
            break;
    }
    if (diffuseLobe) {
        wO = generateDiffuseBounce(noiseSample.z, noiseSample.w, n);
        bsdf = MRT_INV_PI * dot(wO, n); pdf = bsdf;
    } else {
        wO = reflect(wI, n);
        bsdf = dot(wO, n); pdf = 1.0f;
    }
    return wO;
}
__device__ __forceinline__ float lightTriangleSamplePDF(float trianglePdf, float triangleArea, f3 source,       // :181-190
                                                        const Vtx& sample, f3& directionToLight) {
    directionToLight = sample.v - source;
    float distanceToLight = length(directionToLight);
    directionToLight = normalize(directionToLight);
    float LdotD = -dot(directionToLight, sample.n);
    float validDirection = float(distanceToLight >= MRT_DISTANCE_EPSILON) * float(LdotD >= MRT_ANGLE_EPSILON);
    return validDirection * trianglePdf * triangleSamplePDF(triangleArea, LdotD, distanceToLight);
}

// ---- the four kernels of Shaders.metal as device functions on register state ------------------------------
__device__ __forceinline__ float4 load_noise(const float4* tile, uint32_t index) { return __ldg(tile + index); }

// rayGenerator, Shaders.metal:75-103.  t = 0 is hard-wired (:84): side = (1,0,0), up = (0,1,0), view = (0,0,-1).
__device__ __forceinline__ void ray_generator(PathState& st, const float4* noise, uint32_t x, uint32_t y, uint32_t W, uint32_t H) {
    float aspect = __fdiv_rn(float(H), float(W));
    float4 n = load_noise(noise, (x % MRT_NOISE_DIMENSIONS) + (y % MRT_NOISE_DIMENSIONS) * MRT_NOISE_DIMENSIONS);
    float dudvx = __fdiv_rn(n.x * 2.0f - 1.0f, float(W - 1));
    float dudvy = __fdiv_rn(n.y * 2.0f - 1.0f, float(H - 1));
    float ncx = __fdiv_rn(float(2 * x), float(W - 1)) - 1.0f;
    float ncy = __fdiv_rn(float(2 * y), float(H - 1)) - 1.0f;
    st.origin = mk3(0.0f, 1.0f, 2.35f);                                          // up - view * 2.35 (:96)
    st.direction = normalize(mk3(dudvx + ncx, dudvy + ncy * aspect, -1.0f));     // :97
    st.pdf = 1.0f; st.prevDiffuse = 0.0f; st.bounce = 0.0f; st.ior = 1.00029f;   // :99
    st.throughput = mk3(1.0f, 1.0f, 1.0f);
    st.radiance = mk3(0.0f, 0.0f, 0.0f);
}

// intersectionHandler, Shaders.metal:105-212, for a ray that is known to have hit (distance >= DISTANCE_EPSILON;
// the early-out of :122-126 is the caller's).  `noise` = tile of slot (frame + bounce)%3 (Renderer.mm:538).
// Returns true iff the light-sampling ray is live (validLightTriangle, :164,173); `sh` is written only inside
// the NEE branch, exactly like the buffer fields at :171-175.
// EXT = true compiles the optics extension in (alias light selection, refraction, dispersion — chosen by S.extFlags);
// the EXT = false instantiation is the reference's behaviour and the only one launched while S.extFlags == 0.
template <bool EXT = false>
__device__ __forceinline__ bool intersection_handler(const DeviceScene& S, const HitRec& hit, PathState& st, ShadowRay& sh,
                                                     bool& neeBranch, const float4* noise, uint32_t frameIndex,
                                                     uint32_t maxPathLength, uint32_t x, uint32_t y) {
    ShadeTri ref = load_shade_tri(S, hit.tri);
    Mat material = load_material(S, ref.material);
    uint32_t bounce = uint32_t(st.bounce);                                                      // :132
    f3 wI = st.direction;
    uint32_t noiseIndex = ((x + bounce + frameIndex / 3) % MRT_NOISE_DIMENSIONS) +              // :135-136
                          ((y + bounce + frameIndex / 5) % MRT_NOISE_DIMENSIONS) * MRT_NOISE_DIMENSIONS;
    float4 noiseSample = load_noise(noise, noiseIndex);
    Vtx hitVertex = interpolate_uv(ref.t0, ref.t1, ref.t2, hit.u, hit.v);                       // :140
    bool shadowLive = false;
    neeBranch = bounce + 1 < maxPathLength;

    if (neeBranch) {                                                                            // :150-176
        uint32_t selected;
        if (EXT && (S.extFlags & MRT_EXT_ALIAS_LIGHTS)) selected = selectLightAlias(noiseSample.z, S.lightAlias, S.lightCount);
        else selected = selectLightTriangle(noiseSample.z, S.lightCdf, S.lightCount, S.lightGuide, S.lightGuideBuckets);
        Light sel = load_light(S, selected);
        Vtx lightVertex = interpolate(sel.v1, sel.v2, sel.v3, barycentric(noiseSample.w, noiseSample.x));
        f3 directionToLight;
        float lightPdf = lightTriangleSamplePDF(sel.pdf, sel.area, hitVertex.v, lightVertex, directionToLight);
        f2 materialSample = sampleMaterial(material, wI, directionToLight, hitVertex.n, noiseSample);
        float weight = balanceHeuristic(lightPdf, materialSample.y);
        shadowLive = (lightPdf > 0.0f) && (sel.index != hit.tri);
        float scale = __fdiv_rn(weight * materialSample.x, lightPdf);
        f3 e = sel.emissive;
        e = mk3(e.x * material.diffuse.x, e.y * material.diffuse.y, e.z * material.diffuse.z);
        e = mk3(e.x * st.throughput.x, e.y * st.throughput.y, e.z * st.throughput.z);
        sh.throughput = e * scale;
        sh.direction = directionToLight;
        sh.targetIndex = sel.index;
    }

    if (ref.light != MRT_NOT_A_LIGHT) {                                                         // :180-197
        Light sel = load_light(S, ref.light);
        ShadeTri lref = load_shade_tri(S, sel.index);
        Vtx lightVertex = interpolate_uv(lref.t0, lref.t1, lref.t2, hit.u, hit.v);
        f3 directionToLight;
        float mPdf = st.pdf;
        float lPdf = st.prevDiffuse * lightTriangleSamplePDF(sel.pdf, sel.area, st.origin, lightVertex, directionToLight);
        float weight = balanceHeuristic(mPdf, lPdf);
        float k = weight * mPdf;                                                                // sic (:193)
        st.radiance.x += (material.emissive.x * st.throughput.x) * k;
        st.radiance.y += (material.emissive.y * st.throughput.y) * k;
        st.radiance.z += (material.emissive.z * st.throughput.z) * k;
    }

    {                                                                                           // :200-211
        float pdf = 0.0f, bsdf = 0.0f, ior = 0.0f;
        f3 side = hitVertex.n;
        if (EXT && (S.extFlags & MRT_EXT_REFRACTION) && material.type == MRT_MATERIAL_SMOOTH_DIELECTRIC) {
            f3 channelScale;
            st.direction = refractiveBounce(material, S, st.direction, st.ior, st.throughput, hitVertex.n, noiseSample, bsdf, pdf, ior, channelScale);
            st.throughput = mk3(st.throughput.x * channelScale.x, st.throughput.y * channelScale.y, st.throughput.z * channelScale.z);
            if (dot(st.direction, hitVertex.n) < 0.0f) side = -hitVertex.n;      // continue (and sample lights) on the far side
        } else {
            st.direction = generateNextBounce(material, st.direction, st.ior, hitVertex.n, noiseSample, bsdf, pdf, ior);
        }
        st.origin = hitVertex.v + side * MRT_DISTANCE_EPSILON;
        st.pdf = pdf;
        st.prevDiffuse = float(material.type == MRT_MATERIAL_DIFFUSE);
        st.bounce = float(bounce + 1);
        st.ior = ior;
        float r = __fdiv_rn(bsdf, pdf);                                                         // 0/0 when noise.w == 0
        st.throughput.x *= material.diffuse.x * r;
        st.throughput.y *= material.diffuse.y * r;
        st.throughput.z *= material.diffuse.z * r;
    }
    return shadowLive;
}

// accumulateImage, Shaders.metal:233-249, on one pixel.  mode: MRT_ACCUM_*.
__device__ __forceinline__ f3 accumulate_image(f3 color, f3 stored, uint32_t frameIndex, uint32_t mode) {
    if (mode == MRT_ACCUM_MEAN && frameIndex > 0) {
        float factor = __fdiv_rn(float(frameIndex), float(frameIndex + 1));
        color = mk3(mixf(color.x, stored.x, factor), mixf(color.y, stored.y, factor), mixf(color.z, stored.z, factor));
    } else if (mode == MRT_ACCUM_SUM) {
        color = stored + color;
    }
    return color;
}

}  // namespace mrt
