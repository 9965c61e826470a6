// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_abi.h).
//
// orc_kernels.h — scalar C++ restatement of the four compute kernels of renderer/Shaders.metal:75-249.
// Each function processes ONE thread of the reference's 2-D grid (threadId = (x, y), wholeSize = (W, H),
// rayIndex = y*W + x, Shaders.metal:81,117,223,239) on the same buffers the Metal kernels bind.
#pragma once
#include <cmath>
#include "orc_shading.h"

namespace orc {

static inline f4 load_noise(const float* tile, uint32_t index) {
    const float* p = tile + 4 * (size_t)index;
    return f4{p[0], p[1], p[2], p[3]};
}

// Shaders.metal:75-103.  t = 0 is hard-wired (:84) so side = (1,0,0), up = (0,1,0), view = (0,0,-1) exactly.
static inline void rayGenerator(Ray* rays, const float* noise, uint32_t x, uint32_t y, uint32_t W, uint32_t H) {
    uint32_t rayIndex = y * W + x;
    float aspect = float(H) / float(W);
    f4 noiseSample = load_noise(noise, (x % NOISE_DIMENSIONS) + (y % NOISE_DIMENSIONS) * NOISE_DIMENSIONS);
    float dudvx = (noiseSample.x * 2.0f - 1.0f) / float(W - 1);
    float dudvy = (noiseSample.y * 2.0f - 1.0f) / float(H - 1);
    float ncx = float(2 * x) / float(W - 1) - 1.0f;
    float ncy = float(2 * y) / float(H - 1) - 1.0f;
    Ray& r = rays[rayIndex];
    r.origin[0] = 0.0f; r.origin[1] = 1.0f; r.origin[2] = 2.35f;            // up - view * 2.35 (:96)
    f3 dir = normalize(f3{dudvx + ncx, dudvy + ncy * aspect, -1.0f});        // :97 (dudv.y not scaled by aspect)
    st3(r.direction, dir);
    r.maxDistance = INFINITY;
    r.params[0] = 1.0f; r.params[1] = 0.0f; r.params[2] = 0.0f; r.params[3] = 1.00029f;
    spectrum_set(r.throughput, 1.0f);
    spectrum_set(r.radiance, 0.0f);
    // minDistance is never written by any kernel: it keeps the zero the buffer was allocated with
}

// Shaders.metal:105-212.  `noise` is the tile bound at index 8 = slot (frame+i)%3 (Renderer.mm:538).
static inline void intersectionHandler(const Intersection* intersections, const Vertex* vertexBuffer,
                                       const TriangleReference* referenceBuffer, const Material* materialBuffer,
                                       const SharedData& sharedData, const LightTriangle* lightTriangles,
                                       Ray* primaryRays, const float* noise, LightSamplingRay* lightSamplingRays,
                                       uint32_t maxPathLength, uint32_t x, uint32_t y, uint32_t W,
                                       const Extensions& ext = Extensions()) {            // default = the reference
    uint32_t rayIndex = y * W + x;
    lightSamplingRays[rayIndex].maxDistance = -1.0f;                                       // :119
    const Intersection& intersection = intersections[rayIndex];
    if (intersection.distance < DISTANCE_EPSILON) {                                        // :122-126
        primaryRays[rayIndex].maxDistance = -1.0f;
        return;
    }
    Ray& ray = primaryRays[rayIndex];
    const TriangleReference& ref = referenceBuffer[intersection.triangleIndex];
    const Material& material = materialBuffer[ref.materialIndex];

    uint32_t bounce = uint32_t(ray.params[2]);                                             // :132
    f3 wI = ld3(ray.direction);
    uint32_t noiseIndex = ((x + bounce + sharedData.frameIndex / 3) % NOISE_DIMENSIONS) +  // :135-136
                          ((y + bounce + sharedData.frameIndex / 5) % NOISE_DIMENSIONS) * NOISE_DIMENSIONS;
    f4 noiseSample = load_noise(noise, noiseIndex);

    f2 uv{intersection.coordinates[0], intersection.coordinates[1]};
    Vtx hitVertex = interpolate(load_vertex(vertexBuffer[ref.tri[0]]), load_vertex(vertexBuffer[ref.tri[1]]),
                                load_vertex(vertexBuffer[ref.tri[2]]), uv);                // :140

    // light sampling (:150-176)
    if (bounce + 1 < maxPathLength) {
        const LightTriangle& selected = (ext.flags & EXT_ALIAS_LIGHTS)
            ? lightTriangles[selectLightAlias(noiseSample.z, ext.aliasProb, ext.aliasIndex, sharedData.lightTrianglesCount)]
            : selectLightTriangle(noiseSample.z, lightTriangles, (int)sharedData.lightTrianglesCount);
        Vtx lightVertex = interpolate(load_vertex(selected.v1), load_vertex(selected.v2), load_vertex(selected.v3),
                                      barycentric(f2{noiseSample.w, noiseSample.x}));     // noiseSample.wx (:154)
        f3 directionToLight;
        float lightPdf = lightTriangleSamplePDF(selected.pdf, selected.area, hitVertex.v, lightVertex, directionToLight);
        f2 materialSample = sampleMaterial(material, wI, directionToLight, hitVertex.n, noiseSample);
        float materialBsdf = materialSample.x;
        float materialPdf = materialSample.y;
        float weight = balanceHeuristic(lightPdf, materialPdf);
        bool validLightTriangle = (lightPdf > 0.0f) && (selected.index != intersection.triangleIndex);

        Spectrum emittedLight = selected.emissive;
        spectrum_mul_inplace(emittedLight, material.diffuse);
        spectrum_mul_inplace(emittedLight, ray.throughput);
        spectrum_mul_inplace(emittedLight, weight * materialBsdf / lightPdf);

        LightSamplingRay& ls = lightSamplingRays[rayIndex];
        st3(ls.origin, hitVertex.v + hitVertex.n * DISTANCE_EPSILON);
        st3(ls.direction, directionToLight);
        ls.maxDistance = validLightTriangle ? INFINITY : -1.0f;
        ls.targetIndex = selected.index;
        ls.throughput = emittedLight;
    }

    // emitter hit (:180-197); the trailing "* mPdf" is how the reference is written
    if (ref.lightTriangleIndex != uint32_t(-1)) {
        const LightTriangle& selected = lightTriangles[ref.lightTriangleIndex];
        const TriangleReference& lref = referenceBuffer[selected.index];
        Vtx lightVertex = interpolate(load_vertex(vertexBuffer[lref.tri[0]]), load_vertex(vertexBuffer[lref.tri[1]]),
                                      load_vertex(vertexBuffer[lref.tri[2]]), uv);
        f3 directionToLight;
        float mPdf = ray.params[0];
        float lPdf = ray.params[1] * lightTriangleSamplePDF(selected.pdf, selected.area, ld3(ray.origin), lightVertex,
                                                            directionToLight);
        float weight = balanceHeuristic(mPdf, lPdf);
        Spectrum added_radiance = spectrum_mul(material.emissive, ray.throughput);
        spectrum_mul_inplace(added_radiance, weight * mPdf);
        spectrum_add_inplace(ray.radiance, added_radiance);
    }

    // next bounce (:200-211)
    {
        float pdf = 0.0f, bsdf = 0.0f, ior = 0.0f;
        f3 dir, side = hitVertex.n;
        if ((ext.flags & EXT_REFRACTION) && material.materialType == MATERIAL_SMOOTH_DIELECTRIC) {    // extension, see orc_shading.h
            float channelScale[3];
            dir = refractiveBounce(material, ext, ld3(ray.direction), ray.params[3], ray.throughput, hitVertex.n, noiseSample,
                                   bsdf, pdf, ior, channelScale);
            for (int i = 0; i < 3; ++i) ray.throughput.values[i] *= channelScale[i];
            if (dot(dir, hitVertex.n) < 0.0f) side = -hitVertex.n;
            // the light-sampling ray of this hit starts where the path continues from (far side after a transmission)
            if (bounce + 1 < maxPathLength) st3(lightSamplingRays[rayIndex].origin, hitVertex.v + side * DISTANCE_EPSILON);
        } else {
            dir = generateNextBounce(material, ld3(ray.direction), ray.params[3], hitVertex.n, noiseSample, bsdf, pdf, ior);
        }
        st3(ray.direction, dir);
        st3(ray.origin, hitVertex.v + side * DISTANCE_EPSILON);
        ray.maxDistance = INFINITY;
        ray.params[0] = pdf;
        ray.params[1] = float(material.materialType == MATERIAL_DIFFUSE);
        ray.params[2] = float(bounce + 1);
        ray.params[3] = ior;
        Spectrum throughputScale = spectrum_mul(material.diffuse, bsdf / pdf);           // 0/0 when noise.w == 0
        spectrum_mul_inplace(ray.throughput, throughputScale);
    }
}

// Shaders.metal:214-231
static inline void lightSamplingHandler(const Intersection* intersections, Ray* primaryRays,
                                        const LightSamplingRay* lightSamplingRays, uint32_t rayIndex) {
    const Intersection& intersection = intersections[rayIndex];
    if ((intersection.distance >= DISTANCE_EPSILON) &&
        (intersection.triangleIndex == lightSamplingRays[rayIndex].targetIndex)) {
        spectrum_add_inplace(primaryRays[rayIndex].radiance, lightSamplingRays[rayIndex].throughput);
    }
}

// Shaders.metal:233-249.  image = RGBA32F, 4 floats per pixel.  mode 0 = reference running mean,
// 1 = plain sum (sample-sharded multi-GPU; not in the reference), 2 = ACCUMULATE_IMAGE false.
static inline void accumulateImage(float* image, const Ray* sourceRays, uint32_t frameIndex, uint32_t mode,
                                   uint32_t rayIndex) {
    Spectrum color = sourceRays[rayIndex].radiance;
    float* px = image + 4 * (size_t)rayIndex;
    if (mode == 0 && frameIndex > 0) {
        float factor = float(frameIndex) / float(frameIndex + 1);
        for (int i = 0; i < SPECTRUM_SAMPLES; ++i) color.values[i] = mixf(color.values[i], px[i], factor);
    } else if (mode == 1) {
        for (int i = 0; i < SPECTRUM_SAMPLES; ++i) color.values[i] = px[i] + color.values[i];
    }
    px[0] = color.values[0]; px[1] = color.values[1]; px[2] = color.values[2]; px[3] = 1.0f;
}

}  // namespace orc
