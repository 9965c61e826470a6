// kernels.h — host-callable launchers of the sm_100a kernels (implemented in kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "device_shading.cuh"

namespace mrt {

// One entry per frame of a fused batch: frame index + noise-pool tile held by slot 0,1,2 while it renders.
struct FrameEntry { uint32_t frame; uint32_t tile[3]; };
static_assert(sizeof(FrameEntry) == 16, "FrameEntry is one uint4");

struct Grid2D { uint32_t W, H, rowFirst, rowStride; };   // dispatch size + multi-GPU row shard

struct RenderParams {
    DeviceScene S;
    const float4* noisePool;     // MRT_NOISE_DIMENSIONS^2 float4 per tile
    const uint4* frames;         // FrameEntry[nFrames]
    uint32_t nFrames;
    float4* image;
    Grid2D g;
    uint32_t ownedRows;
    uint32_t maxPathLength, accumMode;
    unsigned long long* counters;   // [0] live primary rays, [1] live shadow rays, [2] nodes, [3] triangles
};

// stage kernels — same buffers and bindings as the reference's dispatches (see mrt.h)
void launch_raygen(cudaStream_t s, mrt_ray* rays, const float4* noiseTile, Grid2D g);
void launch_intersect(cudaStream_t s, const DeviceScene& S, const void* rays, uint32_t strideBytes, uint32_t count,
                      mrt_intersection* out, Grid2D g, bool useGrid, unsigned long long* liveCounter,
                      unsigned long long* visitCounters);
void launch_shade(cudaStream_t s, const DeviceScene& S, const mrt_intersection* isects, mrt_ray* rays,
                  mrt_light_sampling_ray* lightRays, const float4* noiseTile, uint32_t frameIndex, uint32_t maxPathLength,
                  Grid2D g);
void launch_shadow_resolve(cudaStream_t s, const mrt_intersection* isects, mrt_ray* rays,
                           const mrt_light_sampling_ray* lightRays, Grid2D g);
void launch_accumulate(cudaStream_t s, float4* image, const mrt_ray* rays, uint32_t frameIndex, uint32_t mode, Grid2D g);
void launch_scale_image(cudaStream_t s, float4* image, float scale, uint32_t count);
// presentation pass (Shaders.metal:33-69); ref may be null (comparison disabled)
void launch_display(cudaStream_t s, const float4* image, const float4* ref, float4* out, uint32_t W, uint32_t H, uint32_t RW,
                    uint32_t RH, uint32_t toneMapping, uint32_t manualSRGB, uint32_t mode, float scale);

// persistent (dynamic-fetch) form of launch_intersect: same arguments and results; `head` = zeroed work counter
void launch_intersect_persistent(cudaStream_t s, const DeviceScene& S, const void* rays, uint32_t strideBytes, uint32_t count,
                                 mrt_intersection* out, Grid2D g, bool useGrid, unsigned long long* liveCounter,
                                 unsigned int* head);

// ---- queue-compacted wavefront: path state as SoA in HBM, one slot per owned pixel ------------------------------
// A wave = k_wf_trace (persistent; extension rays -> nearest hit, light-sampling rays -> visibility) followed by
// k_wf_shade (resolve + intersectionHandler + accumulate/regenerate), which compacts the next wave's two queues
// with warp ballots.  Parity p = wave & 1 selects the queue pair a wave consumes; the other pair is being filled.
struct WfBuffers {
    float4* P0;          // origin.xyz, bsdf pdf            |
    float4* P1;          // direction.xyz, prev-diffuse flag | = the live fields of the reference's Ray
    float4* P2;          // throughput.xyz, bounce           |   (Raytracing.h:54-69)
    float4* P3;          // radiance.xyz, ior                |
    float4* H;           // nearest hit of the extension ray: t, triangle, u, v   (Intersection, Raytracing.h:85-90)
    float4* S0;          // light-sampling ray: direction.xyz, distance of the target along it
    float4* S1;          //                     throughput.xyz, targetIndex      (LightSamplingRay, Raytracing.h:71-83)
    uint32_t* visible;   // written by trace: 1 iff the light-sampling ray's nearest hit is its target (Shaders.metal:225),
                         // 0 if not; written by shade: 2 = light ray queued, 3 = none
    uint32_t* item;      // work item (frame-table entry * pixels + pixel) the slot is tracing; 0xFFFFFFFF none, ..FE retry
    float4* samples;     // radiance of every finished sample of the batch: [frame-table entry][pixel]
    unsigned int* regionNext;   // per region of 256 pixels: next unclaimed item (frame-major inside the region)
    unsigned int* blockRegion;  // per shade block: region it currently draws items from (regions = none)
    uint32_t regions, regionSize;   // regionSize: 32 (one tile, claimed per warp), a multiple of 256, or pixels (one global counter)
    uint32_t pixels;     // owned pixels in 8x4-tiled order, padded: tilesX * tilesY * 32
    uint32_t* queueExt[2];
    uint32_t* queueShadow[2];
    unsigned int* counts;    // [parity][0 ext, 1 shadow]
    unsigned int* heads;     // [parity] work counter of the persistent trace kernel
    const uint32_t* triSlot; // original triangle index -> record in triRecords
    const float4* triRecords;// the triangle records the trace kernel uses (wide or binary tree order)
    uint32_t slots;          // path slots of this group (pointers above are already offset to the group's first slot)
    uint32_t slotBase;       // first slot of the group in the whole slot array (multiple of 256)
    const unsigned int* allCounts; uint32_t groups;   // every group's counts block (8 words each), for the drain check
    unsigned int* waveLog;                            // profile mode: where THIS trace launch notes its two queue sizes (else null)
    // batch control block (device), see MRT_CTL_*: k_wf_shade counts the samples it parks, block 0 of k_wf_trace decides
    // whether the batch is over (nothing queued and `expected` samples parked) and then ends the graph's WHILE loop /
    // tells the polling host loop to stop launching waves
    unsigned int* ctl;
    uint32_t expected;               // samples the batch must finish: owned pixels x frames
    uint32_t waveLimit;              // trace launches after which the loop ends regardless (failure: k_wf_resolve_samples flags it)
    unsigned long long condHandle;   // cudaGraphConditionalHandle of the WHILE node that repeats the wave pair (0 = host loop)
    volatile unsigned int* hostStop; // mapped host word: receives batchSerial when the batch is over (host loop only; else null)
    uint32_t batchSerial;
    // ray ordering (MRT_WF_SORT): shade appends (slot, key) to the staging arrays and counts keys; k_wf_sort scatters the
    // slots into the queues bucket by bucket, so that a trace warp's refills pull rays that start in the same cell of
    // the scene box with the same direction octant
    uint32_t sortBits;               // grid bits per axis (0 = ordering off)
    uint32_t sortMode;               // bit 0: octant-major keys (else cell-major)
    uint32_t* sortSlot[2];           // [0 ext, 1 shadow] unsorted slot ids of the wave being built
    uint32_t* sortKey[2];            //                   their keys
    unsigned int* sortHist;          // [2][buckets] key counts, then [2][buckets] scatter cursors
    float sortLo[3], sortScale[3];   // cell = (origin - lo) * scale, clamped
};
#define MRT_CTL_DONE 0          /* samples parked so far in this batch */
#define MRT_CTL_WAVES 1         /* trace launches of this batch that found rays */
#define MRT_CTL_FINISHED 2      /* 1 once the trace kernel has seen the end of the batch */
#define MRT_CTL_LAUNCHES 3      /* trace launches of this batch, empty ones included (the wave limit counts these) */
#define MRT_CTL_BATCH_WORDS 4   /* words 0..3 are zeroed at the start of every batch */
#define MRT_CTL_GRAPH_TRACES 4  /* trace launches made by graph replays since mrt_reset_stats (kernel_launches accounting) */
#define MRT_CTL_GRAPH_WAVES 5   /* ... of which found rays */
#define MRT_CTL_WORDS 8
__host__ __device__ inline uint32_t wf_sort_buckets(uint32_t bits) { return bits ? 8u << (3u * bits) : 0u; }
void launch_wf_sort(cudaStream_t s, const WfBuffers& b, uint32_t outParity);
void launch_wf_shade(cudaStream_t s, const RenderParams& p, const WfBuffers& b, uint32_t wave, bool first, bool countRays);
void launch_wf_trace(cudaStream_t s, const RenderParams& p, const WfBuffers& b, uint32_t wave, bool countVisits);
void launch_wf_resolve_samples(cudaStream_t s, const RenderParams& p, const WfBuffers& b);
bool trace_uses_wide_bvh();   // MRT_BVH_WIDTH=2 selects the binary tree (A/B runs); default 8
int wf_kernel_info(int* traceRegs, int* traceBlocksPerSM, int* shadeRegs);

// fused megakernel: whole frames, ray state in registers
void launch_render_fused(cudaStream_t s, const RenderParams& p, bool countRays);
int fused_kernel_info(int* regs, int* maxBlocksPerSM);

}  // namespace mrt
