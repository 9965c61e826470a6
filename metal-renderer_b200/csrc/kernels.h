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

// fused megakernel: whole frames, ray state in registers
void launch_render_fused(cudaStream_t s, const RenderParams& p, bool countRays);
int fused_kernel_info(int* regs, int* maxBlocksPerSM);

}  // namespace mrt
