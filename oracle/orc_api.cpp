// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_abi.h).  Not linked into, imported by, or shipped with the
// product library; it is the CPU checker and the timed CPU baseline.
//
// orc_api.cpp — host sequencing of the reference (renderer/Renderer.mm:472-498 updateSharedData,
// 500-585 performRaytracing, 640-657 drawableSizeWillChange) around the kernel restatements in
// orc_kernels.h and the MPS stand-in of orc_bvh.h, behind a small C interface for ctypes.
//
// PARITY STATUS: the reference ships no tests, golden vectors or CPU path for this pipeline, and its
// intersector is closed source, so the END-TO-END oracle is "parity unpinned" (SURVEY.md §8c).  What is
// pinned: struct layouts and the host-side functions of Raytracing.h against the reference header compiled
// here (oracle/ref_probe.cpp -> tests/golden/*.json), analytic known-answer tests per function, and — statistically,
// not bit for bit — the whole pipeline against the Mitsuba images the reference ships as its comparison targets
// (renderer/Media/reference/*.exr -> tests/golden/mitsuba_reference.npz; tests/test_mitsuba_sanity.py: mean radiance
// within 0.3 % at path lengths 2, 3 and 8 at Mitsuba's resolution, 2.7 % at the CPU test's 100x75).
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include "orc_bvh.h"
#include "orc_kernels.h"
#include "orc_display.h"

using namespace orc;

struct orc_ctx {
    uint32_t maxPathLength = 8, accumMode = 0, threads = 1, animateNoise = 1;
    uint32_t W = 0, H = 0, rowFirst = 0, rowStride = 1;
    bool useBvh = true;
    std::vector<Vertex> vertices;
    std::vector<uint32_t> indices;
    std::vector<TriangleReference> references;
    std::vector<Material> materials;
    std::vector<LightTriangle> lights;       // lightCount + 1 (sentinel)
    uint32_t lightCount = 0;
    Bvh bvh;
    std::vector<Ray> rays;
    std::vector<LightSamplingRay> lightRays;
    std::vector<Intersection> intersections;
    std::vector<float> image;
    std::vector<float> noise[NOISE_SLOTS];
    std::atomic<uint64_t> primaryRays{0}, shadowRays{0}, nodesVisited{0}, trisTested{0};
    bool countVisits = false;
    Extensions ext;                          // optics extension, off unless orc_set_extensions is called
    std::vector<float> aliasProb; std::vector<uint32_t> aliasIndex;
};

static void intersect_range(orc_ctx* c, const char* rayBase, size_t stride, Intersection* out, size_t begin, size_t end,
                            uint64_t* live, uint64_t* nv, uint64_t* tt, bool bvh) {
    for (size_t i = begin; i < end; ++i) {
        const float* r = reinterpret_cast<const float*>(rayBase + i * stride);
        float tmin = r[3], tmax = r[7];
        Hit h{-1.0f, 0xFFFFFFFFu, 0.0f, 0.0f};
        if (tmax >= 0.0f) {                      // MPS skips rays with a negative maxDistance
            ++*live;
            f3 o{r[0], r[1], r[2]}, d{r[4], r[5], r[6]};
            h = bvh ? c->bvh.nearest(o, d, tmin, tmax, nv, tt) : c->bvh.nearest_brute(o, d, tmin, tmax);
        }
        out[i].distance = h.t; out[i].triangleIndex = h.tri; out[i].coordinates[0] = h.u; out[i].coordinates[1] = h.v;
    }
}

template <class F>
static void parallel_chunks(uint32_t threads, size_t n, size_t chunk, F&& fn) {
    if (threads <= 1 || n <= chunk) { for (size_t b = 0; b < n; b += chunk) fn(b, std::min(n, b + chunk)); return; }
    std::atomic<size_t> next{0};
    std::vector<std::thread> pool;
    for (uint32_t t = 0; t < threads; ++t)
        pool.emplace_back([&] {
            for (;;) { size_t b = next.fetch_add(chunk); if (b >= n) break; fn(b, std::min(n, b + chunk)); }
        });
    for (auto& th : pool) th.join();
}

static inline bool owns_row(const orc_ctx* c, uint32_t y) { return y % c->rowStride == c->rowFirst; }

// One frame over the pixel range [begin, end): the dispatch order of performRaytracing (Renderer.mm:504-584),
// run chunk-by-chunk instead of buffer-by-buffer (every thread of the reference only touches its own rayIndex,
// so the two orders give identical bits).
static void frame_range(orc_ctx* c, uint32_t frame, const float* const slotTiles[3], size_t begin, size_t end) {
    const uint32_t W = c->W, H = c->H;
    SharedData sd{frame, c->lightCount, 0.0f};
    uint64_t livePrimary = 0, liveShadow = 0, nv = 0, tt = 0;
    uint64_t* pnv = c->countVisits ? &nv : nullptr; uint64_t* ptt = c->countVisits ? &tt : nullptr;
    // pixels of rows this ctx does not own are left untouched (multi-GPU row sharding)
    auto each = [&](auto&& fn) {
        for (size_t i = begin; i < end; ++i) { uint32_t y = (uint32_t)(i / W), x = (uint32_t)(i % W); if (owns_row(c, y)) fn(i, x, y); }
    };
    each([&](size_t, uint32_t x, uint32_t y) { rayGenerator(c->rays.data(), slotTiles[frame % 3], x, y, W, H); });
    for (uint32_t i = 0; i < c->maxPathLength; ++i) {
        each([&](size_t k, uint32_t, uint32_t) {
            intersect_range(c, (const char*)c->rays.data(), sizeof(Ray), c->intersections.data(), k, k + 1, &livePrimary, pnv, ptt, c->useBvh); });
        each([&](size_t, uint32_t x, uint32_t y) {
            intersectionHandler(c->intersections.data(), c->vertices.data(), c->references.data(), c->materials.data(), sd,
                                c->lights.data(), c->rays.data(), slotTiles[(frame + i) % 3], c->lightRays.data(),
                                c->maxPathLength, x, y, W, c->ext); });
        each([&](size_t k, uint32_t, uint32_t) {
            intersect_range(c, (const char*)c->lightRays.data(), sizeof(LightSamplingRay), c->intersections.data(), k, k + 1, &liveShadow, pnv, ptt, c->useBvh); });
        each([&](size_t k, uint32_t, uint32_t) { lightSamplingHandler(c->intersections.data(), c->rays.data(), c->lightRays.data(), (uint32_t)k); });
    }
    each([&](size_t k, uint32_t, uint32_t) { accumulateImage(c->image.data(), c->rays.data(), frame, c->accumMode, (uint32_t)k); });
    c->primaryRays += livePrimary; c->shadowRays += liveShadow; c->nodesVisited += nv; c->trisTested += tt;
}

extern "C" {

orc_ctx* orc_create(uint32_t maxPathLength, uint32_t accumMode, uint32_t threads) {
    orc_ctx* c = new orc_ctx();
    c->maxPathLength = maxPathLength; c->accumMode = accumMode;
    c->threads = threads ? threads : std::max(1u, std::thread::hardware_concurrency());
    for (auto& n : c->noise) n.assign(4 * NOISE_DIMENSIONS * NOISE_DIMENSIONS, 0.0f);
    return c;
}
void orc_destroy(orc_ctx* c) { delete c; }
uint32_t orc_threads(orc_ctx* c) { return c->threads; }
void orc_set_options(orc_ctx* c, uint32_t useBvh, uint32_t countVisits, uint32_t animateNoise) {
    c->useBvh = useBvh != 0; c->countVisits = countVisits != 0; c->animateNoise = animateNoise;
}
void orc_set_accumulate_mode(orc_ctx* c, uint32_t mode) { c->accumMode = mode; }
// optics extension (beyond the reference; orc_shading.h).  The alias table is an input, like the light cdf is.
void orc_set_extensions(orc_ctx* c, uint32_t flags, float dispersion, const float* aliasProb, const uint32_t* aliasIndex, uint32_t n) {
    c->aliasProb.clear(); c->aliasIndex.clear();
    if (aliasProb && aliasIndex) { c->aliasProb.assign(aliasProb, aliasProb + n); c->aliasIndex.assign(aliasIndex, aliasIndex + n); }
    c->ext.flags = flags; c->ext.dispersion = (flags & EXT_DISPERSION) ? dispersion : 0.0f;
    c->ext.aliasProb = c->aliasProb.data(); c->ext.aliasIndex = c->aliasIndex.data();
}
void orc_refractiveBounce(const Material* m, uint32_t flags, float dispersion, const float* dir, float rayIor, const float* throughput3,
                          const float* n, const float* noise4, float* out3, float* bsdfPdfIor3, float* channelScale3) {
    Extensions e; e.flags = flags; e.dispersion = dispersion;
    Spectrum t; for (int i = 0; i < 3; ++i) t.values[i] = throughput3[i];
    float bsdf = 0, pdf = 0, ior = 0;
    st3(out3, refractiveBounce(*m, e, ld3(dir), rayIor, t, ld3(n), f4{noise4[0], noise4[1], noise4[2], noise4[3]}, bsdf, pdf, ior, channelScale3));
    bsdfPdfIor3[0] = bsdf; bsdfPdfIor3[1] = pdf; bsdfPdfIor3[2] = ior;
}
uint32_t orc_selectLightAlias(float xi, const float* prob, const uint32_t* alias, uint32_t n) { return selectLightAlias(xi, prob, alias, n); }
void orc_set_row_shard(orc_ctx* c, uint32_t first, uint32_t stride) { c->rowFirst = first; c->rowStride = stride ? stride : 1; }

// buffer upload + acceleration structure of initRaytracing (Renderer.mm:450-469)
void orc_set_scene(orc_ctx* c, const Vertex* v, uint32_t nv, const uint32_t* idx, uint32_t ntri,
                   const TriangleReference* refs, const Material* mats, uint32_t nmat,
                   const LightTriangle* lights, uint32_t nlights) {
    c->vertices.assign(v, v + nv);
    c->indices.assign(idx, idx + 3 * (size_t)ntri);
    c->references.assign(refs, refs + ntri);
    c->materials.assign(mats, mats + nmat);
    c->lights.assign(lights, lights + nlights + 1);
    c->lightCount = nlights;
    c->bvh.build(c->vertices.data(), c->indices.data(), ntri);
}

// drawableSizeWillChange (Renderer.mm:640-657): zero-filled buffers, accumulation restarts
void orc_resize(orc_ctx* c, uint32_t W, uint32_t H) {
    c->W = W; c->H = H;
    size_t n = (size_t)W * H;
    c->rays.assign(n, Ray{}); c->lightRays.assign(n, LightSamplingRay{}); c->intersections.assign(n, Intersection{});
    c->image.assign(4 * n, 0.0f);
}
void orc_set_noise(orc_ctx* c, uint32_t slot, const float* tile) {
    c->noise[slot % NOISE_SLOTS].assign(tile, tile + 4 * NOISE_DIMENSIONS * NOISE_DIMENSIONS);
}

// ---- single dispatches over the whole grid (test seams at the reference's stage boundaries) --------
void orc_raygen(orc_ctx* c, uint32_t frame) {
    parallel_chunks(c->threads, (size_t)c->W * c->H, 8192, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i) { uint32_t y = (uint32_t)(i / c->W), x = (uint32_t)(i % c->W);
            if (owns_row(c, y)) rayGenerator(c->rays.data(), c->noise[frame % 3].data(), x, y, c->W, c->H); } });
}
void orc_intersect(orc_ctx* c, uint32_t which) {
    const char* base = which == 0 ? (const char*)c->rays.data() : (const char*)c->lightRays.data();
    size_t stride = which == 0 ? sizeof(Ray) : sizeof(LightSamplingRay);
    parallel_chunks(c->threads, (size_t)c->W * c->H, 4096, [&](size_t b, size_t e) {
        uint64_t live = 0, nv = 0, tt = 0;
        for (size_t i = b; i < e; ++i) if (owns_row(c, (uint32_t)(i / c->W)))
            intersect_range(c, base, stride, c->intersections.data(), i, i + 1, &live, c->countVisits ? &nv : nullptr,
                            c->countVisits ? &tt : nullptr, c->useBvh);
        (which == 0 ? c->primaryRays : c->shadowRays) += live; c->nodesVisited += nv; c->trisTested += tt; });
}
void orc_shade(orc_ctx* c, uint32_t frame, uint32_t bounceIteration) {
    SharedData sd{frame, c->lightCount, 0.0f};
    parallel_chunks(c->threads, (size_t)c->W * c->H, 8192, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i) { uint32_t y = (uint32_t)(i / c->W), x = (uint32_t)(i % c->W);
            if (owns_row(c, y)) intersectionHandler(c->intersections.data(), c->vertices.data(), c->references.data(),
                c->materials.data(), sd, c->lights.data(), c->rays.data(), c->noise[(frame + bounceIteration) % 3].data(),
                c->lightRays.data(), c->maxPathLength, x, y, c->W, c->ext); } });
}
void orc_shadow_resolve(orc_ctx* c) {
    parallel_chunks(c->threads, (size_t)c->W * c->H, 8192, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i) if (owns_row(c, (uint32_t)(i / c->W)))
            lightSamplingHandler(c->intersections.data(), c->rays.data(), c->lightRays.data(), (uint32_t)i); });
}
void orc_accumulate(orc_ctx* c, uint32_t frame) {
    parallel_chunks(c->threads, (size_t)c->W * c->H, 8192, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i) if (owns_row(c, (uint32_t)(i / c->W)))
            accumulateImage(c->image.data(), c->rays.data(), frame, c->accumMode, (uint32_t)i); });
}

// performRaytracing for one frame with the current slot contents (Renderer.mm:500-585)
void orc_render_frame(orc_ctx* c, uint32_t frame) {
    const float* tiles[3] = {c->noise[0].data(), c->noise[1].data(), c->noise[2].data()};
    parallel_chunks(c->threads, (size_t)c->W * c->H, 2048, [&](size_t b, size_t e) { frame_range(c, frame, tiles, b, e); });
}

// Frame loop of drawInMTKView (Renderer.mm:587-638) without the display: for each frame
// first, first+stride, ...: updateSharedData refreshes slot f%3 from noiseTiles (tile k <-> frame first+k;
// intermediate frames of a strided loop still refresh their slots, as a full run would), then performRaytracing.
// Returns wall seconds spent in the frame loop.
double orc_render_frames(orc_ctx* c, uint32_t first, uint32_t nFrames, uint32_t stride, const float* noiseTiles) {
    const size_t T = 4 * NOISE_DIMENSIONS * NOISE_DIMENSIONS;
    auto t0 = std::chrono::steady_clock::now();
    if (stride == 0) stride = 1;
    uint32_t last = first + (nFrames ? (nFrames - 1) * stride : 0);
    for (uint32_t f = first, k = 0; nFrames && f <= last; ++f, ++k) {
        if (noiseTiles && c->animateNoise) c->noise[f % 3].assign(noiseTiles + k * T, noiseTiles + (k + 1) * T);
        if ((f - first) % stride == 0) orc_render_frame(c, f);
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void orc_read_buffer(orc_ctx* c, uint32_t which, void* dst, size_t bytes) {
    const void* src = which == 0 ? (const void*)c->rays.data() : which == 1 ? (const void*)c->lightRays.data()
                    : which == 2 ? (const void*)c->intersections.data() : (const void*)c->image.data();
    std::memcpy(dst, src, bytes);
}
void orc_write_buffer(orc_ctx* c, uint32_t which, const void* src, size_t bytes) {
    void* dst = which == 0 ? (void*)c->rays.data() : which == 1 ? (void*)c->lightRays.data()
              : which == 2 ? (void*)c->intersections.data() : (void*)c->image.data();
    std::memcpy(dst, src, bytes);
}
void orc_read_image(orc_ctx* c, float* rgba) { std::memcpy(rgba, c->image.data(), c->image.size() * sizeof(float)); }
void orc_get_counts(orc_ctx* c, uint64_t* out4) {
    out4[0] = c->primaryRays; out4[1] = c->shadowRays; out4[2] = c->nodesVisited; out4[3] = c->trisTested;
}
void orc_reset_counts(orc_ctx* c) { c->primaryRays = 0; c->shadowRays = 0; c->nodesVisited = 0; c->trisTested = 0; }

// one MPSRayIntersector encode over host rays (Renderer.mm:519-523); mode 0 = brute force, 1 = BVH
void orc_intersect_host(orc_ctx* c, const void* rays, size_t stride, uint32_t count, Intersection* out, uint32_t mode) {
    parallel_chunks(c->threads, count, 1024, [&](size_t b, size_t e) {
        uint64_t live = 0;
        intersect_range(c, (const char*)rays, stride, out, b, e, &live, nullptr, nullptr, mode != 0); });
}

// ---- per-function known-answer seams --------------------------------------------------------------
float orc_fresnel(const float* n, const float* i, float etaOut, float etaIn) { return fresnel(ld3(n), ld3(i), etaOut, etaIn); }
float orc_balanceHeuristic(float f, float g) { return balanceHeuristic(f, g); }
float orc_triangleSamplePDF(float area, float cosTheta, float dist) { return triangleSamplePDF(area, cosTheta, dist); }
void orc_barycentric(float x, float y, float* out3) { st3(out3, barycentric(f2{x, y})); }
void orc_buildOrthonormalBasis(const float* n, float* u, float* v) { f3 uu, vv; buildOrthonormalBasis(ld3(n), uu, vv); st3(u, uu); st3(v, vv); }
void orc_generateDiffuseBounce(float sx, float sy, const float* n, float* out3) { st3(out3, generateDiffuseBounce(f2{sx, sy}, ld3(n))); }
void orc_sincos(float x, float* s, float* c) { sincos_pinned(x, s, c); }
void orc_reflect(const float* i, const float* n, float* out3) { st3(out3, reflect(ld3(i), ld3(n))); }
void orc_interpolate(const Vertex* t0, const Vertex* t1, const Vertex* t2, float u, float v, Vertex* out) {
    Vtx r = interpolate(load_vertex(*t0), load_vertex(*t1), load_vertex(*t2), f2{u, v}); st3(out->v, r.v); st3(out->n, r.n);
}
uint32_t orc_selectLightTriangle(float xi, const LightTriangle* lights, int count) {
    return (uint32_t)(&selectLightTriangle(xi, lights, count) - lights);
}
void orc_sampleMaterial(const Material* m, const float* wI, const float* wO, const float* n, const float* noise4, float* out2) {
    f2 r = sampleMaterial(*m, ld3(wI), ld3(wO), ld3(n), f4{noise4[0], noise4[1], noise4[2], noise4[3]}); out2[0] = r.x; out2[1] = r.y;
}
void orc_generateNextBounce(const Material* m, const float* dir, float rayIor, const float* n, const float* noise4,
                            float* outDir, float* bsdfPdfIor) {
    float bsdf, pdf, ior;
    f3 w = generateNextBounce(*m, ld3(dir), rayIor, ld3(n), f4{noise4[0], noise4[1], noise4[2], noise4[3]}, bsdf, pdf, ior);
    st3(outDir, w); bsdfPdfIor[0] = bsdf; bsdfPdfIor[1] = pdf; bsdfPdfIor[2] = ior;
}
float orc_lightTriangleSamplePDF(float triPdf, float area, const float* source, const Vertex* sample, float* dir3) {
    f3 d; float r = lightTriangleSamplePDF(triPdf, area, ld3(source), load_vertex(*sample), d); st3(dir3, d); return r;
}
void orc_triangle_test(const float* o, const float* d, float tmin, float tmax, const float* v0, const float* v1,
                       const float* v2, float* tuv, uint32_t* hit) {
    Hit best{tmax, 0xFFFFFFFFu, 0.0f, 0.0f};
    triangle_test(ld3(o), ld3(d), tmin, tmax, ld3(v0), ld3(v1), ld3(v2), 0, best);
    *hit = best.tri == 0; tuv[0] = best.t; tuv[1] = best.u; tuv[2] = best.v;
}
// presentation pass (Shaders.metal:33-69): tone mapping / sRGB / comparison against a reference image
void orc_display(const float* image, uint32_t W, uint32_t H, uint32_t toneMapping, uint32_t manualSRGB, uint32_t mode, float scale,
                 const float* ref, uint32_t RW, uint32_t RH, float* out) {
    DisplayParams p{toneMapping, manualSRGB, mode, scale};
    display(image, W, H, p, ref, RW, RH, out);
}
float orc_toSRGB(float v) { return toSRGB(v); }

uint32_t orc_sizeof(uint32_t which) {
    switch (which) { case 0: return sizeof(Ray); case 1: return sizeof(LightSamplingRay); case 2: return sizeof(Intersection);
        case 3: return sizeof(Vertex); case 4: return sizeof(Material); case 5: return sizeof(TriangleReference);
        case 6: return sizeof(LightTriangle); case 7: return sizeof(SharedData); case 8: return sizeof(Spectrum); }
    return 0;
}

}  // extern "C"
