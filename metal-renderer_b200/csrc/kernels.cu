// kernels.cu — sm_100a kernels of the per-pixel pipeline raygen -> intersect -> shade/NEE -> shadow intersect ->
// resolve -> accumulate (reference: renderer/Shaders.metal:75-249 sequenced by renderer/Renderer.mm:500-585).
//
// Two forms of the same arithmetic (device_shading.cuh / device_trace.cuh):
//   * stage kernels that keep the reference's buffers and stage boundaries (drop-in + test seams), and
//   * the fused megakernel: one launch renders whole frames with the path state in registers, lanes that finish
//     a sample start the next one at once, and primary and shadow rays share one traversal site.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false (see device_math.cuh for why).
#include <cstdlib>
#include "kernels.h"
#include "device_trace8.cuh"

namespace mrt {

__device__ __forceinline__ bool owns_row(const Grid2D& g, uint32_t y) { return y % g.rowStride == g.rowFirst; }

// ---- rayGenerator (Shaders.metal:75-103) -----------------------------------------------------------------
__global__ void __launch_bounds__(256) k_raygen(mrt_ray* __restrict__ rays, const float4* __restrict__ noise, Grid2D g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.W * g.H) return;
    uint32_t x = i % g.W, y = i / g.W;
    if (!owns_row(g, y)) return;
    PathState st;
    ray_generator(st, noise, x, y, g.W, g.H);
    float4* r = reinterpret_cast<float4*>(rays + i);
    float minDistance = rays[i].minDistance;               // never written by the reference; keep what is there
    r[0] = make_float4(st.origin.x, st.origin.y, st.origin.z, minDistance);
    r[1] = make_float4(st.direction.x, st.direction.y, st.direction.z, INFINITY);
    rays[i].throughput.values[0] = 1.0f; rays[i].throughput.values[1] = 1.0f; rays[i].throughput.values[2] = 1.0f;
    rays[i].radiance.values[0] = 0.0f; rays[i].radiance.values[1] = 0.0f; rays[i].radiance.values[2] = 0.0f;
    r[4] = make_float4(st.pdf, st.prevDiffuse, st.bounce, st.ior);
}

// ---- MPS nearest-hit stand-in over a strided ray buffer (Renderer.mm:519-523, 545-553) -------------------
template <bool COUNT>
__global__ void __launch_bounds__(256) k_intersect(DeviceScene S, const char* __restrict__ rays, uint32_t stride, uint32_t count,
                                                   mrt_intersection* __restrict__ out, Grid2D g, bool useGrid, bool vec,
                                                   unsigned long long* liveCounter, unsigned long long* visitCounters) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long nodesV = 0, trisT = 0;
    uint32_t live = 0;
    if (i < count && (!useGrid || owns_row(g, i / g.W))) {
        const char* p = rays + (size_t)i * stride;
        float4 a, b;
        if (vec) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 16); }
        else {
            const float* f = reinterpret_cast<const float*>(p);
            a = make_float4(f[0], f[1], f[2], f[3]); b = make_float4(f[4], f[5], f[6], f[7]);
        }
        HitRec h; h.t = -1.0f; h.tri = MRT_NO_HIT; h.u = 0.0f; h.v = 0.0f;
        if (b.w >= 0.0f) {                                 // rays with a negative maxDistance are skipped
            live = 1;
            h = trace_nearest<COUNT>(S.nodes, S.tris, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), a.w, b.w, &nodesV, &trisT);
        }
        *reinterpret_cast<float4*>(out + i) = make_float4(h.t, __uint_as_float(h.tri), h.u, h.v);
    }
    if (liveCounter) {
        uint32_t n = __reduce_add_sync(0xffffffffu, live);
        if ((threadIdx.x & 31) == 0 && n) atomicAdd(liveCounter, (unsigned long long)n);
    }
    if (COUNT) {
        for (int o = 16; o; o >>= 1) { nodesV += __shfl_xor_sync(0xffffffffu, nodesV, o); trisT += __shfl_xor_sync(0xffffffffu, trisT, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(visitCounters, nodesV); atomicAdd(visitCounters + 1, trisT); }
    }
}

// ---- intersectionHandler (Shaders.metal:105-212) ----------------------------------------------------------
template <bool EXT>
__global__ void __launch_bounds__(256) k_shade(DeviceScene S, const mrt_intersection* __restrict__ isects, mrt_ray* __restrict__ rays,
                                               mrt_light_sampling_ray* __restrict__ lightRays, const float4* __restrict__ noise,
                                               uint32_t frameIndex, uint32_t maxPathLength, Grid2D g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.W * g.H) return;
    uint32_t x = i % g.W, y = i / g.W;
    if (!owns_row(g, y)) return;
    lightRays[i].maxDistance = -1.0f;                                                   // :119
    float4 is = *reinterpret_cast<const float4*>(isects + i);
    if (is.x < MRT_DISTANCE_EPSILON) { rays[i].maxDistance = -1.0f; return; }            // :122-126
    float4* r = reinterpret_cast<float4*>(rays + i);
    float4 r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3], r4 = r[4];
    PathState st;
    st.origin = mk3(r0.x, r0.y, r0.z); st.direction = mk3(r1.x, r1.y, r1.z);
    st.throughput = mk3(r2.x, r2.y, r2.z); st.radiance = mk3(r2.w, r3.x, r3.y);
    st.pdf = r4.x; st.prevDiffuse = r4.y; st.bounce = r4.z; st.ior = r4.w;
    HitRec hit; hit.t = is.x; hit.tri = __float_as_uint(is.y); hit.u = is.z; hit.v = is.w;
    ShadowRay sh; bool nee;
    bool live = intersection_handler<EXT>(S, hit, st, sh, nee, noise, frameIndex, maxPathLength, x, y);
    if (nee) {                                                                          // :171-175
        float4* l = reinterpret_cast<float4*>(lightRays + i);
        float lmin = lightRays[i].minDistance;
        l[0] = make_float4(st.origin.x, st.origin.y, st.origin.z, lmin);
        l[1] = make_float4(sh.direction.x, sh.direction.y, sh.direction.z, live ? INFINITY : -1.0f);
        l[2] = make_float4(sh.throughput.x, sh.throughput.y, sh.throughput.z, __uint_as_float(sh.targetIndex));
    }
    r[0] = make_float4(st.origin.x, st.origin.y, st.origin.z, r0.w);
    r[1] = make_float4(st.direction.x, st.direction.y, st.direction.z, INFINITY);
    r[2] = make_float4(st.throughput.x, st.throughput.y, st.throughput.z, st.radiance.x);
    r[3] = make_float4(st.radiance.y, st.radiance.z, r3.z, r3.w);
    r[4] = make_float4(st.pdf, st.prevDiffuse, st.bounce, st.ior);
}

// ---- lightSamplingHandler (Shaders.metal:214-231) -----------------------------------------------------------
__global__ void __launch_bounds__(256) k_shadow_resolve(const mrt_intersection* __restrict__ isects, mrt_ray* __restrict__ rays,
                                                        const mrt_light_sampling_ray* __restrict__ lightRays, Grid2D g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.W * g.H || !owns_row(g, i / g.W)) return;
    float4 is = *reinterpret_cast<const float4*>(isects + i);
    float4 l2 = reinterpret_cast<const float4*>(lightRays + i)[2];
    if (is.x >= MRT_DISTANCE_EPSILON && __float_as_uint(is.y) == __float_as_uint(l2.w)) {
        rays[i].radiance.values[0] += l2.x; rays[i].radiance.values[1] += l2.y; rays[i].radiance.values[2] += l2.z;
    }
}

// ---- accumulateImage (Shaders.metal:233-249) ----------------------------------------------------------------
__global__ void __launch_bounds__(256) k_accumulate(float4* __restrict__ image, const mrt_ray* __restrict__ rays, uint32_t frameIndex,
                                                    uint32_t mode, Grid2D g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.W * g.H || !owns_row(g, i / g.W)) return;
    f3 color = mk3(rays[i].radiance.values[0], rays[i].radiance.values[1], rays[i].radiance.values[2]);
    float4 stored = image[i];
    color = accumulate_image(color, mk3(stored.x, stored.y, stored.z), frameIndex, mode);
    image[i] = make_float4(color.x, color.y, color.z, 1.0f);
}

__global__ void __launch_bounds__(256) k_scale_image(float4* __restrict__ image, float scale, uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 c = image[i];
    image[i] = make_float4(c.x * scale, c.y * scale, c.z * scale, 1.0f);
}

// ---- presentation pass: blitFragment (Shaders.metal:33-69) as a compute kernel over a headless RGBA32F buffer ----
// One thread per output pixel; the output has the image's size, so the nearest sample of the image is the pixel
// itself and the reference (any size) is read at its nearest texel.  toSRGB = Raytracing.h:130-135.
__device__ __forceinline__ float to_srgb(float v) {
    if (v <= 0.0f) return 0.0f;
    if (v >= 1.0f) return 1.0f;
    return (v < 0.0031308f) ? (12.92f * v) : (1.055f * powf(v, 1.0f / 2.4f) - 0.055f);
}
__device__ __forceinline__ uint32_t nearest_texel(uint32_t i, uint32_t n, uint32_t m) {
    const float u = __fdiv_rn(float(i) + 0.5f, float(n));
    const uint32_t t = uint32_t(u * float(m));
    return t < m ? t : m - 1u;
}
__global__ void __launch_bounds__(256) k_display(const float4* __restrict__ image, const float4* __restrict__ ref,
                                                 float4* __restrict__ out, uint32_t W, uint32_t H, uint32_t RW, uint32_t RH,
                                                 uint32_t toneMapping, uint32_t manualSRGB, uint32_t mode, float scale) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    const uint32_t x = i % W, y = i / W;
    float4 c = image[i];
    if (toneMapping) c = make_float4(1.0f - expf(-c.x), 1.0f - expf(-c.y), 1.0f - expf(-c.z), 1.0f - expf(-c.w));
    if (manualSRGB) { c.x = to_srgb(c.x); c.y = to_srgb(c.y); c.z = to_srgb(c.z); }
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ref == nullptr) mode = MRT_COMPARE_DISABLED;
    if (mode != MRT_COMPARE_DISABLED) r = ref[(size_t)nearest_texel(y, H, RH) * RW + nearest_texel(x, W, RW)];
    float4 o = c;
    if (mode == MRT_COMPARE_ABSOLUTE_VALUE)
        o = make_float4(fabsf(c.x - r.x) * scale, fabsf(c.y - r.y) * scale, fabsf(c.z - r.z) * scale, fabsf(c.w - r.w) * scale);
    else if (mode == MRT_COMPARE_REF_TO_COLOR)
        o = make_float4(fmaxf(0.0f, r.x - c.x) * scale, fmaxf(0.0f, r.y - c.y) * scale, fmaxf(0.0f, r.z - c.z) * scale, fmaxf(0.0f, r.w - c.w) * scale);
    else if (mode == MRT_COMPARE_COLOR_TO_REF)
        o = make_float4(fmaxf(0.0f, c.x - r.x) * scale, fmaxf(0.0f, c.y - r.y) * scale, fmaxf(0.0f, c.z - r.z) * scale, fmaxf(0.0f, c.w - r.w) * scale);
    else if (mode == MRT_COMPARE_LUMINANCE) {
        const float third = 1.0f / 3.0f;
        const float lc = dot(mk3(c.x, c.y, c.z), mk3(third, third, third)), lr = dot(mk3(r.x, r.y, r.z), mk3(third, third, third));
        o = make_float4(fmaxf(0.0f, lc - lr) * scale, fmaxf(0.0f, lr - lc) * scale, 0.0f * scale, 1.0f * scale);
    }
    out[i] = o;
}

// ---- fused megakernel ----------------------------------------------------------------------------------------
// One thread owns one pixel for every frame of the batch; a warp owns an 8x4 pixel tile.  Per lane a small state
// machine runs GEN -> (PRIMARY -> [SHADOW])* -> GEN ...: a lane whose sample ends (path died or reached
// maxPathLength) accumulates it and generates the next frame's camera ray in the same iteration, so every
// traversal sees a full set of live rays, and primary and shadow rays go through ONE traversal site.  The
// sequence of float operations per pixel is exactly the stage sequence of Renderer.mm:504-584.
enum { PH_GEN = 0, PH_PRIMARY = 1, PH_SHADOW = 2 };

// WIDE: the traversal site is the wavefront pipeline's own core (device_trace8.cuh: 8-wide quantised tree, vote
// scheduling) run over the warp's 32 rays without refill, so that megakernel and wavefront differ in the SCHEDULE only.
// The 32 "work items" of a call are the warp's lanes; a warp-private counter in shared memory stands in for the queue head.
struct FusedIO {
    f3 o, d; bool want; HitRec out;
    __device__ __forceinline__ bool load(uint32_t, TraceJob& job) {
        if (!want) return false;
        job.o = o; job.d = d; job.tmin = 0.0f; job.tmax = INFINITY; job.target = MRT_NO_HIT; job.tag = 0;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t, uint32_t, const HitRec& h) { out = h; }
};

template <bool COUNT, int MINB, bool WIDE>
__global__ void __launch_bounds__(256, MINB) k_render_fused(RenderParams P, uint32_t leafMin, uint32_t magic) {
    __shared__ __align__(16) uint8_t orderLut[WIDE ? MRT_ORDER_LUT_BYTES : 4];
    __shared__ unsigned int warpHead[8];
    if (WIDE) bvh8_fill_order_lut(orderLut);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t tilesX = (P.g.W + 7u) >> 3;
    const uint32_t tx = warp % tilesX, ty = warp / tilesX;
    const uint32_t x = tx * 8u + (lane & 7u);
    const uint32_t row = ty * 4u + (lane >> 3);
    const uint32_t y = P.g.rowFirst + row * P.g.rowStride;
    bool done = !(x < P.g.W && row < P.ownedRows);
    const uint32_t pix = done ? 0u : y * P.g.W + x;

    f3 color = mk3(0.0f, 0.0f, 0.0f);
    if (!done) { float4 c = P.image[pix]; color = mk3(c.x, c.y, c.z); }
    PathState st; ShadowRay sh;
    st.origin = st.direction = st.throughput = st.radiance = mk3(0.0f, 0.0f, 0.0f);
    st.pdf = st.prevDiffuse = st.bounce = st.ior = 0.0f;
    sh.direction = sh.throughput = mk3(0.0f, 0.0f, 0.0f); sh.targetIndex = 0;
    uint32_t phase = PH_GEN, k = 0, frame = 0, tile0 = 0, tile1 = 0, tile2 = 0;
    uint32_t nPrimary = 0, nShadow = 0;
    unsigned long long dummyA = 0, dummyB = 0;
    const size_t tileStride = MRT_NOISE_DIMENSIONS * MRT_NOISE_DIMENSIONS;

    for (;;) {
        if (!done && phase == PH_GEN) {
            if (k == P.nFrames) {
                P.image[pix] = make_float4(color.x, color.y, color.z, 1.0f);
                done = true;
            } else {
                uint4 e = __ldg(P.frames + k);
                ++k;
                frame = e.x; tile0 = e.y; tile1 = e.z; tile2 = e.w;
                uint32_t slot = frame % 3u;
                uint32_t t = slot == 0 ? tile0 : (slot == 1 ? tile1 : tile2);
                ray_generator(st, P.noisePool + t * tileStride, x, y, P.g.W, P.g.H);      // Renderer.mm:508-512
                phase = PH_PRIMARY;
            }
        }
        if (__all_sync(0xffffffffu, done)) break;
        const f3 dir = phase == PH_PRIMARY ? st.direction : sh.direction;
        HitRec hit;
        if (WIDE) {                                       // warp-collective: finished lanes come along without a ray
            FusedIO io; io.o = st.origin; io.d = dir; io.want = !done;
            io.out.t = -1.0f; io.out.tri = MRT_NO_HIT; io.out.u = io.out.v = 0.0f;
            if (lane == 0) warpHead[threadIdx.x >> 5] = 0u;
            __syncwarp();
            trace_persistent8(P.S.nodes8, P.S.tris8, 32u, &warpHead[threadIdx.x >> 5], 32u, leafMin, magic, orderLut, io);
            __syncwarp();
            hit = io.out;
            if (done) continue;
        } else {
            if (done) continue;
            hit = trace_nearest<false>(P.S.nodes, P.S.tris, st.origin, dir, 0.0f, INFINITY, &dummyA, &dummyB);
        }

        bool finished = false;
        if (phase == PH_PRIMARY) {
            if (COUNT) ++nPrimary;
            if (hit.t < MRT_DISTANCE_EPSILON) {                                        // Shaders.metal:122-126
                finished = true;
            } else {
                uint32_t bounce = uint32_t(st.bounce);
                uint32_t slot = (frame + bounce) % 3u;                                  // Renderer.mm:538
                uint32_t t = slot == 0 ? tile0 : (slot == 1 ? tile1 : tile2);
                bool nee;
                bool live = intersection_handler(P.S, hit, st, sh, nee, P.noisePool + t * tileStride, frame,
                                                 P.maxPathLength, x, y);
                if (live) phase = PH_SHADOW;
                else finished = uint32_t(st.bounce) >= P.maxPathLength;
            }
        } else {
            if (COUNT) ++nShadow;
            if (hit.t >= MRT_DISTANCE_EPSILON && hit.tri == sh.targetIndex)             // Shaders.metal:225-229
                st.radiance = st.radiance + sh.throughput;
            phase = PH_PRIMARY;
            finished = uint32_t(st.bounce) >= P.maxPathLength;
        }
        if (finished) {
            color = accumulate_image(st.radiance, color, frame, P.accumMode);          // Shaders.metal:239-248
            phase = PH_GEN;
        }
    }
    if (COUNT) {
        uint32_t a = __reduce_add_sync(0xffffffffu, nPrimary), b = __reduce_add_sync(0xffffffffu, nShadow);
        if (lane == 0) { atomicAdd(P.counters, (unsigned long long)a); atomicAdd(P.counters + 1, (unsigned long long)b); }
    }
}

// ---- launchers -----------------------------------------------------------------------------------------------
static inline uint32_t blocks_for(uint32_t n) { return (n + 255u) / 256u; }

void launch_raygen(cudaStream_t s, mrt_ray* rays, const float4* noiseTile, Grid2D g) {
    k_raygen<<<blocks_for(g.W * g.H), 256, 0, s>>>(rays, noiseTile, g);
}
void launch_intersect(cudaStream_t s, const DeviceScene& S, const void* rays, uint32_t strideBytes, uint32_t count,
                      mrt_intersection* out, Grid2D g, bool useGrid, unsigned long long* liveCounter,
                      unsigned long long* visitCounters) {
    if (count == 0) return;
    bool vec = (strideBytes % 16u == 0) && ((uintptr_t)rays % 16u == 0);
    const char* r = static_cast<const char*>(rays);
    if (visitCounters) k_intersect<true><<<blocks_for(count), 256, 0, s>>>(S, r, strideBytes, count, out, g, useGrid, vec, liveCounter, visitCounters);
    else k_intersect<false><<<blocks_for(count), 256, 0, s>>>(S, r, strideBytes, count, out, g, useGrid, vec, liveCounter, nullptr);
}
void launch_shade(cudaStream_t s, const DeviceScene& S, const mrt_intersection* isects, mrt_ray* rays,
                  mrt_light_sampling_ray* lightRays, const float4* noiseTile, uint32_t frameIndex, uint32_t maxPathLength,
                  Grid2D g) {
    if (S.extFlags) k_shade<true><<<blocks_for(g.W * g.H), 256, 0, s>>>(S, isects, rays, lightRays, noiseTile, frameIndex, maxPathLength, g);
    else k_shade<false><<<blocks_for(g.W * g.H), 256, 0, s>>>(S, isects, rays, lightRays, noiseTile, frameIndex, maxPathLength, g);
}
void launch_shadow_resolve(cudaStream_t s, const mrt_intersection* isects, mrt_ray* rays,
                           const mrt_light_sampling_ray* lightRays, Grid2D g) {
    k_shadow_resolve<<<blocks_for(g.W * g.H), 256, 0, s>>>(isects, rays, lightRays, g);
}
void launch_accumulate(cudaStream_t s, float4* image, const mrt_ray* rays, uint32_t frameIndex, uint32_t mode, Grid2D g) {
    k_accumulate<<<blocks_for(g.W * g.H), 256, 0, s>>>(image, rays, frameIndex, mode, g);
}
void launch_scale_image(cudaStream_t s, float4* image, float scale, uint32_t count) {
    k_scale_image<<<blocks_for(count), 256, 0, s>>>(image, scale, count);
}
void launch_display(cudaStream_t s, const float4* image, const float4* ref, float4* out, uint32_t W, uint32_t H, uint32_t RW,
                    uint32_t RH, uint32_t toneMapping, uint32_t manualSRGB, uint32_t mode, float scale) {
    k_display<<<blocks_for(W * H), 256, 0, s>>>(image, ref, out, W, H, RW, RH, toneMapping, manualSRGB, mode, scale);
}
// The production library carries ONE megakernel: the per-lane binary-tree loop at 2 blocks per SM, which shares neither
// tree nor traversal code with the wavefront pipeline — that independence is what the "schedules agree" tests lean on.
// -DMRT_EXPERIMENTS adds MRT_FUSED_BVH=8 (the megakernel on the wide traversal core, the like-for-like comparison with the
// wavefront pipeline of DESIGN.md 3) and MRT_FUSED_MINBLOCKS.
#ifdef MRT_EXPERIMENTS
static int fused_bvh_width() {
    static const int w = [] { const char* e = getenv("MRT_FUSED_BVH"); return (e && atoi(e) == 8) ? 8 : 2; }();
    return w;
}
#endif
void launch_render_fused(cudaStream_t s, const RenderParams& p, bool countRays) {
    uint32_t tilesX = (p.g.W + 7u) >> 3, tilesY = (p.ownedRows + 3u) >> 2;
    uint32_t warps = tilesX * tilesY;
    if (warps == 0 || p.nFrames == 0) return;
    uint32_t blocks = (warps + 7u) / 8u;
    const uint32_t magic = 0x47000000u;
#define MRT_LAUNCH_FUSED(MB, WD) do { if (countRays) k_render_fused<true, MB, WD><<<blocks, 256, 0, s>>>(p, leafMin, magic); \
                                      else k_render_fused<false, MB, WD><<<blocks, 256, 0, s>>>(p, leafMin, magic); } while (0)
#ifdef MRT_EXPERIMENTS
    // blocks per SM: 3 on the wide core (80 registers with some spill beat 127 registers at 2 blocks, pass AB: +17 % / +11 %)
    static const int minb = [] { const char* e = getenv("MRT_FUSED_MINBLOCKS"); int v = e ? atoi(e) : (fused_bvh_width() == 8 ? 3 : 2); return v; }();
    // vote tuning of the wide traversal core, packed as in launch_wf_trace (wavefront.cu): 10 lanes / ratio 7
    static const uint32_t leafMin = [] {
        const char* a = getenv("MRT_TRACE_LEAFMIN"); const char* b = getenv("MRT_TRACE_TRIRATIO");
        int l = a ? atoi(a) : 10, r = b ? atoi(b) : 7;
        return (uint32_t)(l < 1 ? 1 : (l > 32 ? 32 : l)) | ((uint32_t)(r < 0 ? 0 : (r > 255 ? 255 : r)) << 8);
    }();
    if (fused_bvh_width() == 8) {
        if (minb >= 4) MRT_LAUNCH_FUSED(4, true); else if (minb == 3) MRT_LAUNCH_FUSED(3, true); else MRT_LAUNCH_FUSED(2, true);
        return;
    }
    if (minb >= 4) { MRT_LAUNCH_FUSED(4, false); return; }
    if (minb == 3) { MRT_LAUNCH_FUSED(3, false); return; }
#else
    const uint32_t leafMin = 10u | (7u << 8);
#endif
    MRT_LAUNCH_FUSED(2, false);
#undef MRT_LAUNCH_FUSED
}
int fused_kernel_info(int* regs, int* maxBlocksPerSM) {
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, k_render_fused<true, 2, false>);
    if (e != cudaSuccess) return (int)e;
    *regs = a.numRegs;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(maxBlocksPerSM, k_render_fused<true, 2, false>, 256, 0);
}

}  // namespace mrt
