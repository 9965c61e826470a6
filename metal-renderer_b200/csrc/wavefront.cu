// wavefront.cu — queue-compacted wavefront form of the per-pixel pipeline, and the persistent intersector.
//
// Same arithmetic as the stage kernels / megakernel of kernels.cu (device_shading.cuh, device_trace.cuh), hence the
// same bits; what changes is the schedule.  The reference runs 2 + 4*L full-size dispatches per frame over AoS
// buffers with dead rays left in place (renderer/Renderer.mm:504-584).  Here:
//   * every owned pixel has one path SLOT whose state (the live fields of Ray / LightSamplingRay / Intersection,
//     Raytracing.h:54-90) lives in SoA float4 arrays in HBM; a slot that finishes a sample accumulates it
//     (Shaders.metal:239-248) and starts the next frame's sample at once, so the extension queue stays full;
//   * k_wf_shade = lightSamplingHandler (Shaders.metal:214-231) of the previous bounce, then intersectionHandler
//     (:105-212) of this one, then accumulate + rayGenerator (:75-103) when the sample ends; it appends the live
//     extension / light-sampling rays of the next wave to two queues with warp ballots + one atomic per block;
//   * k_wf_trace is persistent (one resident grid): warps pull queue entries from a global counter and refill
//     idle lanes while the others keep traversing (device_trace.cuh: trace_persistent).  Light-sampling rays are
//     first tested against their target triangle (in k_wf_shade) and then only need "does anything beat
//     (tTarget, target)?" — the exact predicate of Shaders.metal:225 — so they stop at the first occluder.
#include <cstdlib>
#include "kernels.h"
#include "device_trace.cuh"
#include "device_trace8.cuh"
#ifdef MRT_EXPERIMENTS
#include "device_trace8_experiments.cuh"
#endif

namespace mrt {

__device__ __forceinline__ bool wf_owns_row(const Grid2D& g, uint32_t y) { return y % g.rowStride == g.rowFirst; }

// ---- IO policy over a strided AoS ray buffer (the MPS drop-in, Renderer.mm:519-523 / 545-553) -------------------
struct StridedIO {
    const char* rays; uint32_t stride; mrt_intersection* out; Grid2D g; bool useGrid, vec;
    uint32_t live;
    __device__ __forceinline__ bool load(uint32_t i, TraceJob& job) {
        if (useGrid && !wf_owns_row(g, i / g.W)) return false;
        const char* p = rays + (size_t)i * stride;
        float4 a, b;
        if (vec) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 16); }
        else {
            const float* f = reinterpret_cast<const float*>(p);
            a = make_float4(f[0], f[1], f[2], f[3]); b = make_float4(f[4], f[5], f[6], f[7]);
        }
        if (!(b.w >= 0.0f)) {                              // skipped ray: reports a miss
            *reinterpret_cast<float4*>(out + i) = make_float4(-1.0f, __uint_as_float(MRT_NO_HIT), 0.0f, 0.0f);
            return false;
        }
        ++live;
        job.o = mk3(a.x, a.y, a.z); job.d = mk3(b.x, b.y, b.z); job.tmin = a.w; job.tmax = b.w;
        job.target = MRT_NO_HIT; job.tag = i;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, uint32_t, const HitRec& h) {
        *reinterpret_cast<float4*>(out + i) = make_float4(h.t, __uint_as_float(h.tri), h.u, h.v);
    }
};

template <int MINB, int WIDE>
__global__ void __launch_bounds__(128, MINB) k_intersect_persistent(DeviceScene S, const char* __restrict__ rays, uint32_t stride,
                                                                    uint32_t count, mrt_intersection* __restrict__ out, Grid2D g,
                                                                    bool useGrid, bool vec, unsigned long long* liveCounter,
                                                                    unsigned int* head, uint32_t refillBelow, uint32_t leafMin, uint32_t magic) {
    StridedIO io; io.rays = rays; io.stride = stride; io.out = out; io.g = g; io.useGrid = useGrid; io.vec = vec; io.live = 0;
    __shared__ __align__(16) uint8_t orderLut[WIDE ? MRT_ORDER_LUT_BYTES : 4];
    if (WIDE) bvh8_fill_order_lut(orderLut);
    if (WIDE) trace_persistent8(S.nodes8, S.tris8, count, head, refillBelow, leafMin, magic, orderLut, io);
    else trace_persistent(S.nodes, S.tris, count, head, refillBelow, leafMin, io);      // -DMRT_EXPERIMENTS builds only
    if (liveCounter) {
        uint32_t n = __reduce_add_sync(0xffffffffu, io.live);
        if ((threadIdx.x & 31) == 0 && n) atomicAdd(liveCounter, (unsigned long long)n);
    }
}

// ---- IO policy over the wavefront queues ------------------------------------------------------------------------
struct WfIO {
    const float4 *P0, *P1, *S0, *S1;
    float4* H; uint32_t* visible;
    const uint32_t *qe, *qs;
    uint32_t nExt;
    __device__ __forceinline__ bool load(uint32_t i, TraceJob& job) {
        if (i < nExt) {                                    // extension ray: nearest hit, [0, inf)  (Renderer.mm:519-523)
            uint32_t slot = qe[i];
            float4 a = P0[slot], b = P1[slot];
            job.o = mk3(a.x, a.y, a.z); job.d = mk3(b.x, b.y, b.z); job.tmin = 0.0f; job.tmax = INFINITY;
            job.target = MRT_NO_HIT; job.tag = slot;
        } else {                                           // light-sampling ray, bounded by its target (Renderer.mm:545-553)
            uint32_t slot = qs[i - nExt];
            float4 a = P0[slot], b = S0[slot], c = S1[slot];
            job.o = mk3(a.x, a.y, a.z); job.d = mk3(b.x, b.y, b.z); job.tmin = 0.0f; job.tmax = b.w;
            job.target = __float_as_uint(c.w); job.tag = slot;
        }
        return true;
    }
    __device__ __forceinline__ void store(uint32_t slot, uint32_t target, const HitRec& h) {
        if (target == MRT_NO_HIT) H[slot] = make_float4(h.t, __uint_as_float(h.tri), h.u, h.v);
        else visible[slot] = (h.tri == target) ? 1u : 0u;
    }
};

template <int MINB, int WIDE, bool COUNT = false>
__global__ void __launch_bounds__(128, MINB) k_wf_trace(DeviceScene S, WfBuffers b, uint32_t parity, uint32_t refillBelow,
                                                        uint32_t leafMin, uint32_t magic, unsigned long long* counters) {
    const uint32_t nExt = b.counts[2 * parity], nShadow = b.counts[2 * parity + 1];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // the other queue pair was consumed by the previous wave's trace; clear it for the shade that follows us
        b.counts[2 * (parity ^ 1u)] = 0; b.counts[2 * (parity ^ 1u) + 1] = 0; b.heads[parity ^ 1u] = 0;
        if (counters && nExt) atomicAdd(counters, (unsigned long long)nExt);        // live extension rays traversed
        if (b.waveLog) { b.waveLog[0] = nExt; b.waveLog[1] = nShadow; }             // mrt_set_profile: per-wave queue sizes
        // batch control: the batch is over when nothing is queued and every sample has been parked (k_wf_shade counts
        // them).  Under the graph's WHILE node that ends the loop; the polling host loop stops launching waves.
        const unsigned int work = nExt + nShadow ? 1u : 0u;
        b.ctl[MRT_CTL_WAVES] += work;
        const unsigned int launches = ++b.ctl[MRT_CTL_LAUNCHES];
        const bool finished = !work && b.ctl[MRT_CTL_DONE] >= b.expected;
        if (finished) b.ctl[MRT_CTL_FINISHED] = 1u;
        if (b.condHandle) {
            b.ctl[MRT_CTL_GRAPH_TRACES] += 1u; b.ctl[MRT_CTL_GRAPH_WAVES] += work;
            if (finished || launches >= b.waveLimit) cudaGraphSetConditional(b.condHandle, 0u);
        }
        if (finished && b.hostStop) { *b.hostStop = b.batchSerial; __threadfence_system(); }
    }
    if (b.sortBits) {       // key counts and scatter cursors of the wave k_wf_sort has just placed: zero them for the next shade
        const uint32_t n = 4u * wf_sort_buckets(b.sortBits);
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) b.sortHist[i] = 0u;
    }
    WfIO io; io.P0 = b.P0; io.P1 = b.P1; io.S0 = b.S0; io.S1 = b.S1; io.H = b.H; io.visible = b.visible;
    io.qe = b.queueExt[parity]; io.qs = b.queueShadow[parity]; io.nExt = nExt;
    __shared__ __align__(16) uint8_t orderLut[WIDE ? MRT_ORDER_LUT_BYTES : 4];
    if (WIDE) bvh8_fill_order_lut(orderLut);
#ifdef MRT_EXPERIMENTS
    __shared__ TqWarp tq[WIDE == 3 ? 4 : 1];
    if (WIDE == 3) trace_persistent8q<WfIO, COUNT>(S.nodes8, S.tris8, nExt + nShadow, b.heads + parity, refillBelow, leafMin, magic,
                                                   orderLut, &tq[threadIdx.x >> 5], io, counters ? counters + 2 : nullptr);
    else if (WIDE == 2) trace_persistent8d<WfIO, COUNT>(S.nodes8, S.tris8, nExt + nShadow, b.heads + parity, refillBelow, leafMin, magic,
                                                   orderLut, io, counters ? counters + 2 : nullptr);
    else
#endif
    if (WIDE) trace_persistent8<WfIO, COUNT>(S.nodes8, S.tris8, nExt + nShadow, b.heads + parity, refillBelow, leafMin, magic,
                                             orderLut, io, counters ? counters + 2 : nullptr);
    else trace_persistent(S.nodes, S.tris, nExt + nShadow, b.heads + parity, refillBelow, leafMin, io);
}

// ---- k_wf_shade: resolve -> intersectionHandler -> sample done? -> next work item, queue compaction ---------------
// Work item w = k * pixels + pix: sample of frame-table entry k at owned pixel pix (32 consecutive pix = one 8x4 tile,
// so primary rays of neighbouring items stay coherent).  A slot that finishes its sample writes the radiance to the
// per-sample buffer and takes the NEXT unclaimed item, whatever pixel it belongs to: slots never idle while work is
// left (with one slot tied to one pixel, pixels whose paths always run to the length limit kept a thin tail of waves
// alive for L * frames waves — pass G: 1024 launches averaging 77 % of a full wave).  k_wf_resolve_samples then folds
// the samples into the image in frame order, which is what keeps the running mean of Shaders.metal:241-247 bit-exact.
#define MRT_ITEM_NONE  0xFFFFFFFFu      /* slot has nothing to do (and will not get anything: work exhausted) */
#define MRT_ITEM_RETRY 0xFFFFFFFEu      /* slot drew a padding pixel: fetch again next wave */

#define WF_PROBES 8
__device__ __forceinline__ uint32_t wf_region_pixels(uint32_t r, uint32_t regionSize, uint32_t pixels) {
    const uint32_t at = r * regionSize;
    return pixels - at < regionSize ? pixels - at : regionSize;
}

__device__ __forceinline__ bool wf_pixel_of(uint32_t pix, const RenderParams& P, uint32_t& x, uint32_t& y) {
    const uint32_t tilesX = (P.g.W + 7u) >> 3;
    const uint32_t tile = pix >> 5, l = pix & 31u;
    x = (tile % tilesX) * 8u + (l & 7u);
    const uint32_t row = (tile / tilesX) * 4u + (l >> 3);
    y = P.g.rowFirst + row * P.g.rowStride;
    return x < P.g.W && row < P.ownedRows;
}

// Ordering key of a ray: cell of its origin in a 2^bits-per-axis grid over the scene box (Morton order) and the octant of
// its direction.  Rays of one bucket start next to each other and descend the wide tree in the same child order.
__device__ __forceinline__ uint32_t wf_spread3(uint32_t x) {          // 0000abcd -> 00a00b00c00d
    x = (x | (x << 4)) & 0x0C3u;
    return (x | (x << 2)) & 0x249u;
}
__device__ __forceinline__ uint32_t wf_sort_key(const WfBuffers& b, f3 o, f3 d) {
    const float top = float((1u << b.sortBits) - 1u);
    const uint32_t cx = (uint32_t)fminf(fmaxf((o.x - b.sortLo[0]) * b.sortScale[0], 0.0f), top);
    const uint32_t cy = (uint32_t)fminf(fmaxf((o.y - b.sortLo[1]) * b.sortScale[1], 0.0f), top);
    const uint32_t cz = (uint32_t)fminf(fmaxf((o.z - b.sortLo[2]) * b.sortScale[2], 0.0f), top);
    const uint32_t cell = wf_spread3(cx) | (wf_spread3(cy) << 1) | (wf_spread3(cz) << 2);
    const uint32_t oct = (__float_as_uint(d.x) >> 31) | ((__float_as_uint(d.y) >> 31) << 1) | ((__float_as_uint(d.z) >> 31) << 2);
    return (b.sortMode & 1u) ? ((oct << (3u * b.sortBits)) | cell) : ((cell << 3) | oct);
}

template <bool COUNT, int MINB, bool SORT, bool EXT = false>
__global__ void __launch_bounds__(256, MINB) k_wf_shade(RenderParams P, WfBuffers b, uint32_t outParity, uint32_t first, uint32_t seqHost) {
    __shared__ uint32_t sE[8], sS[8], sN[8], sW[8], sD[8], sBase[5];
    const uint32_t seq = seqHost + b.ctl[MRT_CTL_LAUNCHES];        // varies the probe sequence from wave to wave (graph replays included)
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const bool valid = slot < b.slots;
    const size_t tileStride = MRT_NOISE_DIMENSIONS * MRT_NOISE_DIMENSIONS;
    bool pushE = false, pushS = false, need = false, parked = false;
    uint32_t liveShadow = 0, keyE = 0, keyS = 0;
    uint32_t item = MRT_ITEM_NONE;
    PathState st;
    st.origin = st.direction = st.throughput = st.radiance = mk3(0.f, 0.f, 0.f); st.pdf = st.prevDiffuse = st.bounce = st.ior = 0.f;

    if (valid) {
        item = first ? MRT_ITEM_RETRY : b.item[slot];
        if (item == MRT_ITEM_RETRY) need = true;
        else if (item != MRT_ITEM_NONE) {
            const uint32_t k = item / b.pixels, pix = item - k * b.pixels;
            uint32_t x, y;
            wf_pixel_of(pix, P, x, y);
            float4 p0 = b.P0[slot], p1 = b.P1[slot], p2 = b.P2[slot], p3 = b.P3[slot];
            st.origin = mk3(p0.x, p0.y, p0.z); st.pdf = p0.w;
            st.direction = mk3(p1.x, p1.y, p1.z); st.prevDiffuse = p1.w;
            st.throughput = mk3(p2.x, p2.y, p2.z); st.bounce = p2.w;
            st.radiance = mk3(p3.x, p3.y, p3.z); st.ior = p3.w;
            const uint4 e = __ldg(P.frames + k);
            if (b.visible[slot] == 1u) {                                               // Shaders.metal:225-229
                float4 s1 = b.S1[slot]; st.radiance = st.radiance + mk3(s1.x, s1.y, s1.z);
            }
            const float4 h = b.H[slot];
            bool finished;
            if (h.x < MRT_DISTANCE_EPSILON) {                                          // Shaders.metal:122-126
                finished = true;
            } else {
                HitRec hit; hit.t = h.x; hit.tri = __float_as_uint(h.y); hit.u = h.z; hit.v = h.w;
                const uint32_t s = (e.x + uint32_t(st.bounce)) % 3u;                   // Renderer.mm:538
                const uint32_t t = s == 0 ? e.y : (s == 1 ? e.z : e.w);
                ShadowRay sh; bool nee;
                const bool live = intersection_handler<EXT>(P.S, hit, st, sh, nee, P.noisePool + t * tileStride, e.x,
                                                       P.maxPathLength, x, y);
                if (live) {
                    liveShadow = 1;
                    // the light-sampling ray can only count if it reaches its target at t >= DISTANCE_EPSILON
                    HitRec tg; tg.t = INFINITY; tg.tri = MRT_NO_HIT; tg.u = tg.v = 0.0f;
                    triangle_test(b.triRecords, __ldg(b.triSlot + sh.targetIndex), st.origin, sh.direction, 0.0f, INFINITY, tg);
                    if (tg.tri == sh.targetIndex && tg.t >= MRT_DISTANCE_EPSILON) {
                        pushS = true;
                        b.S0[slot] = make_float4(sh.direction.x, sh.direction.y, sh.direction.z, tg.t);
                        b.S1[slot] = make_float4(sh.throughput.x, sh.throughput.y, sh.throughput.z, __uint_as_float(sh.targetIndex));
                        if (SORT) keyS = wf_sort_key(b, st.origin, sh.direction);
                    }
                }
                finished = uint32_t(st.bounce) >= P.maxPathLength;
            }
            if (finished) {                                // sample complete: park its radiance, ask for new work
                b.samples[item] = make_float4(st.radiance.x, st.radiance.y, st.radiance.z, 1.0f);
                need = true; pushS = false; parked = true;
            } else {
                pushE = true;
            }
        }
    }

    // ---- claim work items from the claimer's current region: one atomic per warp (regions of one 8x4 tile) or per
    //      block (larger regions); a claimer whose region ran dry adopts a random region that still has items ----
    const unsigned mW = __ballot_sync(0xffffffffu, need);
    const uint32_t R = b.regions, nf = P.nFrames, rs = b.regionSize;
    uint32_t jBase = 0, jLim = 0, jCur = 0, rank = __popc(mW & lt);
    auto claim = [&](uint32_t claimer, uint32_t own, uint32_t a) {
        uint32_t cur = first ? own : b.blockRegion[claimer], base = 0, lim = 0;
        if (cur < R) {
            lim = wf_region_pixels(cur, rs, b.pixels) * nf;
            base = atomicAdd(b.regionNext + cur, a);
        }
        if (base >= lim) {
            cur = R; base = lim = 0;
            uint32_t h = (claimer + 1u) * 0x9E3779B1u + seq * 0x85EBCA77u;
            uint32_t cand[WF_PROBES], left[WF_PROBES];
#pragma unroll
            for (int t = 0; t < WF_PROBES; ++t) {
                h = h * 0x2C1B3C6Du + 0x297A2D39u;
                cand[t] = (uint32_t)(((unsigned long long)(h ^ (h >> 15)) * R) >> 32);
                left[t] = *(volatile unsigned int*)(b.regionNext + cand[t]);
            }
#pragma unroll
            for (int t = 0; t < WF_PROBES; ++t) {
                const uint32_t l = wf_region_pixels(cand[t], rs, b.pixels) * nf;
                if (cur == R && left[t] < l) {
                    const uint32_t got = atomicAdd(b.regionNext + cand[t], a);
                    if (got < l) { cur = cand[t]; base = got; lim = l; }
                }
            }
        }
        b.blockRegion[claimer] = cur;
        jBase = base; jLim = lim; jCur = cur;
    };
    if (rs == 32u) {
        if (mW) {
            if (lane == 0) claim((b.slotBase + slot) >> 5, ((b.slotBase + slot) >> 5) % R, __popc(mW));
            jBase = __shfl_sync(0xffffffffu, jBase, 0); jLim = __shfl_sync(0xffffffffu, jLim, 0); jCur = __shfl_sync(0xffffffffu, jCur, 0);
        }
    } else {
        if (lane == 0) sW[warp] = __popc(mW);
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t a = 0;
            for (int w = 0; w < 8; ++w) { uint32_t n = sW[w]; sW[w] = a; a += n; }
            if (a) claim(blockIdx.x + b.slotBase / 256u, (uint32_t)((((unsigned long long)blockIdx.x * 256u + b.slotBase) / rs) % R), a);
            sBase[2] = jBase; sBase[3] = jLim; sBase[4] = jCur;
        }
        __syncthreads();
        jBase = sBase[2]; jLim = sBase[3]; jCur = sBase[4]; rank += sW[warp];
    }
    if (need) {
        const uint32_t j = jBase + rank;
        item = MRT_ITEM_RETRY;                             // nothing this wave (region ran dry, or a padding pixel)
        if (j < jLim) {
            const uint32_t rp = wf_region_pixels(jCur, rs, b.pixels);
            const uint32_t k = j / rp, pix = jCur * rs + (j - k * rp);
            uint32_t x, y;
            if (wf_pixel_of(pix, P, x, y)) {
                const uint4 e = __ldg(P.frames + k);
                const uint32_t s = e.x % 3u;
                const uint32_t t = s == 0 ? e.y : (s == 1 ? e.z : e.w);
                ray_generator(st, P.noisePool + t * tileStride, x, y, P.g.W, P.g.H);   // Renderer.mm:508-512
                pushE = true;
                item = k * b.pixels + pix;
            }
        }
    }
    if (valid) {
        if (pushE) {
            b.P0[slot] = make_float4(st.origin.x, st.origin.y, st.origin.z, st.pdf);
            b.P1[slot] = make_float4(st.direction.x, st.direction.y, st.direction.z, st.prevDiffuse);
            b.P2[slot] = make_float4(st.throughput.x, st.throughput.y, st.throughput.z, st.bounce);
            b.P3[slot] = make_float4(st.radiance.x, st.radiance.y, st.radiance.z, st.ior);
            if (SORT) keyE = wf_sort_key(b, st.origin, st.direction);
        }
        b.item[slot] = item;
        b.visible[slot] = pushS ? 2u : 3u;                 // 2 = light ray in flight (trace writes 0/1), 3 = none
    }

    // ---- append to the next wave's queues: warp ballots, one atomic per block and queue ----
    const unsigned mE = __ballot_sync(0xffffffffu, pushE), mS = __ballot_sync(0xffffffffu, pushS);
    const unsigned mD = __ballot_sync(0xffffffffu, parked);
    const uint32_t nLive = COUNT ? __reduce_add_sync(0xffffffffu, liveShadow) : 0u;
    if (lane == 0) { sE[warp] = __popc(mE); sS[warp] = __popc(mS); sN[warp] = nLive; sD[warp] = __popc(mD); }
    __syncthreads();
    // three warps each take one counter, so that the block waits for ONE global atomic round trip instead of two or three
    if (threadIdx.x == 0) {
        uint32_t aE = 0;
        for (int w = 0; w < 8; ++w) { uint32_t e = sE[w]; sE[w] = aE; aE += e; }
        sBase[0] = aE ? atomicAdd(b.counts + 2 * outParity, aE) : 0u;
    } else if (threadIdx.x == 32) {
        uint32_t aS = 0;
        for (int w = 0; w < 8; ++w) { uint32_t q = sS[w]; sS[w] = aS; aS += q; }
        sBase[1] = aS ? atomicAdd(b.counts + 2 * outParity + 1, aS) : 0u;
    } else if (threadIdx.x == 64) {
        uint32_t aD = 0;
        for (int w = 0; w < 8; ++w) aD += sD[w];
        if (aD) atomicAdd(b.ctl + MRT_CTL_DONE, aD);                                   // samples finished (ends the batch)
    } else if (COUNT && threadIdx.x == 96) {
        uint32_t aN = 0;
        for (int w = 0; w < 8; ++w) aN += sN[w];
        if (aN) atomicAdd(P.counters + 1, (unsigned long long)aN);                     // live light-sampling rays
    }
    __syncthreads();
    if (SORT) {            // staging arrays + key counts; k_wf_sort places the slots in the queues bucket by bucket
        const uint32_t B = wf_sort_buckets(b.sortBits);
        if (pushE) {
            const uint32_t at = sBase[0] + sE[warp] + __popc(mE & lt);
            b.sortSlot[0][at] = slot; b.sortKey[0][at] = keyE; atomicAdd(b.sortHist + keyE, 1u);
        }
        if (pushS) {
            const uint32_t at = sBase[1] + sS[warp] + __popc(mS & lt);
            b.sortSlot[1][at] = slot; b.sortKey[1][at] = keyS; atomicAdd(b.sortHist + B + keyS, 1u);
        }
    } else {
        if (pushE) b.queueExt[outParity][sBase[0] + sE[warp] + __popc(mE & lt)] = slot;
        if (pushS) b.queueShadow[outParity][sBase[1] + sS[warp] + __popc(mS & lt)] = slot;
    }
}

#ifdef MRT_EXPERIMENTS
// Counting sort of one wave's two queues by ray key (MRT_WF_SORT).  Every block scans the key counts itself (they are
// complete: k_wf_shade ran before us), then places its share of the staged slots: offset of the bucket + a position
// from the bucket's cursor, taken once per warp and key (__match_any_sync), so that a warp's rays of one key stay
// together.  The order inside a bucket is whatever the atomics give — the image does not depend on it.
__global__ void __launch_bounds__(256) k_wf_sort(WfBuffers b, uint32_t outParity) {
    extern __shared__ uint32_t sOff[];                 // [2][buckets] exclusive offsets
    __shared__ uint32_t sWarp[8];
    const uint32_t B = wf_sort_buckets(b.sortBits);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t per = (B + 255u) / 256u;
    for (uint32_t q = 0; q < 2u; ++q) {
        const unsigned int* hist = b.sortHist + q * B;
        uint32_t sum = 0;
        for (uint32_t i = 0; i < per; ++i) { const uint32_t k = tid * per + i; if (k < B) sum += hist[k]; }
        uint32_t incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, incl, d); if ((int)lane >= d) incl += n; }
        if (lane == 31u) sWarp[warp] = incl;
        __syncthreads();
        uint32_t base = 0;
        for (uint32_t w = 0; w < warp; ++w) base += sWarp[w];
        uint32_t run = base + incl - sum;
        for (uint32_t i = 0; i < per; ++i) { const uint32_t k = tid * per + i; if (k < B) { sOff[q * B + k] = run; run += hist[k]; } }
        __syncthreads();
    }
    for (uint32_t q = 0; q < 2u; ++q) {
        const uint32_t n = b.counts[2 * outParity + q];
        uint32_t* out = q ? b.queueShadow[outParity] : b.queueExt[outParity];
        unsigned int* cursor = b.sortHist + (2u + q) * B;
        for (uint32_t i0 = (blockIdx.x * 256u + (tid & ~31u)); i0 < n; i0 += gridDim.x * 256u) {
            const uint32_t i = i0 + lane;
            const bool ok = i < n;
            const uint32_t key = ok ? b.sortKey[q][i] : 0xFFFFFFFFu;
            const uint32_t slot = ok ? b.sortSlot[q][i] : 0u;
            const unsigned same = __match_any_sync(0xffffffffu, key);
            const int leader = __ffs(same) - 1;
            uint32_t at = 0;
            if (ok && (int)lane == leader) at = atomicAdd(cursor + key, (unsigned int)__popc(same));
            at = __shfl_sync(0xffffffffu, at, leader);
            if (ok) out[sOff[q * B + key] + at + __popc(same & ((1u << lane) - 1u))] = slot;
        }
    }
}
#endif

// accumulateImage (Shaders.metal:233-249) over the batch, in frame order, one thread per owned pixel
__global__ void __launch_bounds__(256) k_wf_resolve_samples(RenderParams P, WfBuffers b) {
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix == 0) {     // every item must have been claimed and both queue pairs drained (api.cu checks this word)
        unsigned int left = 0;
        for (uint32_t g = 0; g < b.groups; ++g) left += b.allCounts[8 * g] + b.allCounts[8 * g + 1] + b.allCounts[8 * g + 2] + b.allCounts[8 * g + 3];
        if (left) b.counts[7] = left;
    }
    if (pix < b.pixels && (pix % b.regionSize) == 0u &&
        b.regionNext[pix / b.regionSize] < wf_region_pixels(pix / b.regionSize, b.regionSize, b.pixels) * P.nFrames) b.counts[7] = 1u;
    uint32_t x, y;
    if (pix >= b.pixels || !wf_pixel_of(pix, P, x, y)) return;
    const uint32_t at = y * P.g.W + x;
    const float4 c = P.image[at];
    f3 color = mk3(c.x, c.y, c.z);
    for (uint32_t k = 0; k < P.nFrames; ++k) {
        const float4 r = b.samples[(size_t)k * b.pixels + pix];
        color = accumulate_image(mk3(r.x, r.y, r.z), color, __ldg(P.frames + k).x, P.accumMode);
    }
    P.image[at] = make_float4(color.x, color.y, color.z, 1.0f);
}

// ---- launchers ---------------------------------------------------------------------------------------------------
// The production library has ONE trace kernel (k_wf_trace<8, 1> and its counting twin), one shade kernel
// (k_wf_shade<., 4, .>) and one persistent intersector (k_intersect_persistent<8, 1>).  Everything else this file can
// launch — other occupancies, the binary tree under the persistent kernel, postponed triangles, the warp-wide triangle
// queue — is a measured experiment and exists only in the -DMRT_EXPERIMENTS build
// (lib/libmrt_b200_exp.so), where the MRT_* environment knobs below select it.
#ifdef MRT_EXPERIMENTS
static int env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
#endif
static int device_sms() {
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n > 0 ? n : 148;
}
struct TraceTuning { int minb, refill, leafMin, wide, triRatio, defer, triThresh, tq; };
static TraceTuning trace_tuning() {
#ifdef MRT_EXPERIMENTS
    static const int wide = env_int("MRT_BVH_WIDTH", 8) == 8 ? 1 : 0;
    static const TraceTuning t = { env_int("MRT_TRACE_MINBLOCKS", wide ? 8 : 10), env_int("MRT_TRACE_REFILL", 24),
                                   env_int("MRT_TRACE_LEAFMIN", 10), wide, env_int("MRT_TRACE_TRIRATIO", 7),
                                   env_int("MRT_TRACE_DEFER", 0), env_int("MRT_TRACE_TRITHRESH", 16), env_int("MRT_TRACE_TQ", 0) };
#else
    // refill below 24 busy lanes; triangle step at >= 10 lanes with triangles or 16 * nT >= 7 * nN (passes P, AW)
    static const TraceTuning t = { 8, 24, 10, 1, 7, 0, 16, 0 };
#endif
    return t;
}
static uint32_t pack_vote(const TraceTuning& t) {
    return (uint32_t)(t.leafMin < 1 ? 1 : (t.leafMin > 32 ? 32 : t.leafMin)) |
           ((uint32_t)(t.triRatio < 0 ? 0 : (t.triRatio > 255 ? 255 : t.triRatio)) << 8) |
           ((uint32_t)(t.triThresh < 1 ? 1 : (t.triThresh > 33 ? 33 : t.triThresh)) << 18);
}

void launch_intersect_persistent(cudaStream_t s, const DeviceScene& S, const void* rays, uint32_t strideBytes, uint32_t count,
                                 mrt_intersection* out, Grid2D g, bool useGrid, unsigned long long* liveCounter,
                                 unsigned int* head) {
    if (count == 0) return;
    const bool vec = (strideBytes % 16u == 0) && ((uintptr_t)rays % 16u == 0);
    const TraceTuning t = trace_tuning();
    const char* r = static_cast<const char*>(rays);
    const uint32_t warpsNeeded = (count + 31u) / 32u;
    const uint32_t refill = (uint32_t)(t.refill < 1 ? 1 : (t.refill > 32 ? 32 : t.refill));
    const uint32_t leafMin = pack_vote(t);
    auto blocks = [&](int perSM) { uint32_t cap = (uint32_t)(device_sms() * perSM); uint32_t need = (warpsNeeded + 3u) / 4u; return need < cap ? need : cap; };
#define MRT_LAUNCH_ISECT(MB, WD) k_intersect_persistent<MB, WD><<<blocks(MB), 128, 0, s>>>(S, r, strideBytes, count, out, g, useGrid, vec, liveCounter, head, refill, leafMin, 0x47000000u)
#ifdef MRT_EXPERIMENTS
    if (t.wide) {
        if (t.minb >= 12) MRT_LAUNCH_ISECT(12, 1); else if (t.minb >= 10) MRT_LAUNCH_ISECT(10, 1);
        else if (t.minb >= 8) MRT_LAUNCH_ISECT(8, 1); else if (t.minb >= 6) MRT_LAUNCH_ISECT(6, 1); else MRT_LAUNCH_ISECT(5, 1);
    } else {
        if (t.minb >= 12) MRT_LAUNCH_ISECT(12, 0); else if (t.minb >= 10) MRT_LAUNCH_ISECT(10, 0);
        else if (t.minb >= 8) MRT_LAUNCH_ISECT(8, 0); else if (t.minb >= 6) MRT_LAUNCH_ISECT(6, 0); else MRT_LAUNCH_ISECT(5, 0);
    }
#else
    MRT_LAUNCH_ISECT(8, 1);
#endif
#undef MRT_LAUNCH_ISECT
}

void launch_wf_trace(cudaStream_t s, const RenderParams& p, const WfBuffers& b, uint32_t wave, bool countVisits) {
    const TraceTuning t = trace_tuning();
    const uint32_t refill = (uint32_t)(t.refill < 1 ? 1 : (t.refill > 32 ? 32 : t.refill));
    const uint32_t leafMin = pack_vote(t);
    const uint32_t parity = wave & 1u;
    const uint32_t maxUseful = (2u * b.slots / 32u + 3u) / 4u;                // both queues full
    auto blocks = [&](int perSM) { uint32_t cap = (uint32_t)(device_sms() * perSM); return maxUseful < cap ? (maxUseful ? maxUseful : 1u) : cap; };
#define MRT_LAUNCH_TRACE(MB, WD) k_wf_trace<MB, WD><<<blocks(MB), 128, 0, s>>>(p.S, b, parity, refill, leafMin, 0x47000000u, p.counters)
    // countVisits (mrt_set_profile bit 1): the counting twin adds node / triangle tests into stats
#ifdef MRT_EXPERIMENTS
    if (t.wide && t.tq && p.S.triangleCount < (1u << 27)) {            // warp-wide triangle queue (device_trace8.cuh); MRT_TRACE_TQ = pairs that trigger a triangle step
        static const int tqWait = env_int("MRT_TRACE_TQWAIT", 8);      // waiting lanes that force a triangle step
        const uint32_t th = (uint32_t)(t.tq < 1 ? 1 : (t.tq > 96 ? 96 : t.tq)) | ((uint32_t)(tqWait < 1 ? 1 : (tqWait > 33 ? 33 : tqWait)) << 8);
        if (countVisits && p.counters) k_wf_trace<8, 3, true><<<blocks(8), 128, 0, s>>>(p.S, b, parity, refill, th, 0x47000000u, p.counters);
        else k_wf_trace<8, 3><<<blocks(8), 128, 0, s>>>(p.S, b, parity, refill, th, 0x47000000u, p.counters);
        return;
    }
    if (t.wide && t.defer) {                                           // postponed triangles (device_trace8.cuh)
        if (countVisits && p.counters) k_wf_trace<8, 2, true><<<blocks(8), 128, 0, s>>>(p.S, b, parity, refill, leafMin, 0x47000000u, p.counters);
        else MRT_LAUNCH_TRACE(8, 2);
        return;
    }
    if (!t.wide) {
        if (t.minb >= 12) MRT_LAUNCH_TRACE(12, 0); else if (t.minb >= 10) MRT_LAUNCH_TRACE(10, 0);
        else if (t.minb >= 8) MRT_LAUNCH_TRACE(8, 0); else if (t.minb >= 6) MRT_LAUNCH_TRACE(6, 0); else MRT_LAUNCH_TRACE(5, 0);
        return;
    }
    if (!(countVisits && p.counters) && t.minb != 8) {
        if (t.minb >= 12) MRT_LAUNCH_TRACE(12, 1); else if (t.minb >= 10) MRT_LAUNCH_TRACE(10, 1);
        else if (t.minb >= 9) MRT_LAUNCH_TRACE(9, 1); else if (t.minb >= 6) MRT_LAUNCH_TRACE(6, 1); else MRT_LAUNCH_TRACE(5, 1);
        return;
    }
#endif
    if (countVisits && p.counters) k_wf_trace<8, 1, true><<<blocks(8), 128, 0, s>>>(p.S, b, parity, refill, leafMin, 0x47000000u, p.counters);
    else MRT_LAUNCH_TRACE(8, 1);
#undef MRT_LAUNCH_TRACE
}

void launch_wf_resolve_samples(cudaStream_t s, const RenderParams& p, const WfBuffers& b) {
    if (b.pixels == 0 || p.nFrames == 0) return;
    k_wf_resolve_samples<<<(b.pixels + 255u) / 256u, 256, 0, s>>>(p, b);
}

void launch_wf_shade(cudaStream_t s, const RenderParams& p, const WfBuffers& b, uint32_t wave, bool first, bool countRays) {
    if (b.slots == 0) return;
    const uint32_t blocks = (b.slots + 255u) / 256u;
    const uint32_t outParity = first ? 0u : ((wave + 1u) & 1u);               // shade(w) fills the queues trace(w+1) consumes
    const uint32_t f = first ? 1u : 0u;
    if (p.S.extFlags) {                                 // optics extension (mrt_set_extensions): its own instantiation, so the
        if (countRays) k_wf_shade<true, 4, false, true><<<blocks, 256, 0, s>>>(p, b, outParity, f, wave);     // default kernel's code is untouched
        else           k_wf_shade<false, 4, false, true><<<blocks, 256, 0, s>>>(p, b, outParity, f, wave);
        return;
    }
    // 4 blocks per SM = 64 registers with a 56-byte spill frame; 3 = 80 registers.  The kernel is latency-bound (ncu pass O:
    // 36 % warps active, long scoreboard), so occupancy wins: +3.6 % per step in pass P
#define MRT_LAUNCH_SHADE_S(MB, SORT) do { if (countRays) k_wf_shade<true, MB, SORT><<<blocks, 256, 0, s>>>(p, b, outParity, f, wave); \
                                          else           k_wf_shade<false, MB, SORT><<<blocks, 256, 0, s>>>(p, b, outParity, f, wave); } while (0)
#ifdef MRT_EXPERIMENTS
#define MRT_LAUNCH_SHADE(MB) do { if (b.sortBits) MRT_LAUNCH_SHADE_S(MB, true); else MRT_LAUNCH_SHADE_S(MB, false); } while (0)
#else
#define MRT_LAUNCH_SHADE(MB) MRT_LAUNCH_SHADE_S(MB, false)      /* ray ordering (MRT_WF_SORT) lost: experiments build only */
#endif
#ifdef MRT_EXPERIMENTS
    static const int minb = env_int("MRT_SHADE_MINBLOCKS", 4);
    if (minb >= 5) { MRT_LAUNCH_SHADE(5); return; }
    if (minb <= 3) { MRT_LAUNCH_SHADE(3); return; }
#endif
    MRT_LAUNCH_SHADE(4);
#undef MRT_LAUNCH_SHADE
#undef MRT_LAUNCH_SHADE_S
}

void launch_wf_sort(cudaStream_t s, const WfBuffers& b, uint32_t outParity) {
#ifdef MRT_EXPERIMENTS
    if (!b.sortBits || b.slots == 0) return;
    const uint32_t need = (b.slots + 255u) / 256u, cap = (uint32_t)device_sms() * 4u;
    k_wf_sort<<<need < cap ? need : cap, 256, 2u * wf_sort_buckets(b.sortBits) * sizeof(uint32_t), s>>>(b, outParity);
#else
    (void)s; (void)b; (void)outParity;
#endif
}

bool trace_uses_wide_bvh() { return trace_tuning().wide != 0; }

int wf_kernel_info(int* traceRegs, int* traceBlocksPerSM, int* shadeRegs) {
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, k_wf_trace<8, 1>);
    if (e != cudaSuccess) return (int)e;
    *traceRegs = a.numRegs;
    e = cudaFuncGetAttributes(&a, k_wf_shade<false, 4, false>);
    if (e != cudaSuccess) return (int)e;
    *shadeRegs = a.numRegs;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(traceBlocksPerSM, k_wf_trace<8, 1>, 128, 0);
}

}  // namespace mrt
