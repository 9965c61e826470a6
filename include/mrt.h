/*
 * mrt.h — C-ABI of the B200 path-tracing core (libmrt_b200.so).
 *
 * The reference has no plugin/FFI interface; the boundary this library replaces is
 *   (1) the Objective-C lifecycle of `Renderer` (reference: renderer/Renderer.h:5-6,
 *       renderer/Renderer.mm:72-87, 255-470, 472-498, 500-585, 640-657), and
 *   (2) the positional buffer bindings of its six GPU dispatches
 *       (renderer/Renderer.mm:508-510, 519-523, 529-539, 545-553, 559-565, 577-579 <->
 *       [[buffer(n)]] in renderer/Shaders.metal:75-79, 105-115, 214-221, 233-237).
 * Every export below cites the reference interface it stands in for.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * MRT_E_* code, with a human-readable message available from mrt_last_error() (thread-local).
 * A ctx owns all device memory and one in-order CUDA stream (the reference uses one
 * MTLCommandQueue, Renderer.mm:153).  Host arrays passed in are borrowed for the duration of the
 * call only.  A ctx is not thread-safe; distinct ctxs are independent.  There is NO CPU fallback:
 * without a usable CUDA device mrt_create fails with MRT_E_CUDA.
 */
#ifndef MRT_H
#define MRT_H

#include "mrt_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MRT_VERSION 104

enum {
    MRT_OK = 0,
    MRT_E_INVALID = -1,   /* bad argument / call order */
    MRT_E_CUDA = -2,      /* CUDA runtime error or no device */
    MRT_E_NOMEM = -3,
    MRT_E_IO = -4,        /* scene file problems */
    MRT_E_STATE = -5      /* scene / size / noise not set yet */
};

/* Traversal / pipeline variant used by mrt_render_frames. */
enum {
    MRT_PIPELINE_AUTO = 0,
    MRT_PIPELINE_MEGAKERNEL = 1,   /* one fused kernel per batch of frames, path state in registers */
    MRT_PIPELINE_WAVEFRONT = 2,    /* queue-compacted waves: persistent dynamic-fetch trace kernel + shade kernel, SoA
                                      path state in HBM, warp-ballot compaction of live rays (AUTO selects this) */
    MRT_PIPELINE_STAGED = 3        /* the reference's own 2 + 4*L dispatches per frame over the ABI buffers */
};

/* How mrt_accumulate / mrt_render_frames fold a frame into the image. */
enum {
    MRT_ACCUM_MEAN = 0,   /* reference behaviour: c = mix(c, stored, f/(f+1)), Shaders.metal:241-247 */
    MRT_ACCUM_SUM = 1,    /* image += radiance (sample-sharded multi-GPU; reduce, then mrt_scale_image) */
    MRT_ACCUM_REPLACE = 2 /* ACCUMULATE_IMAGE == false, Raytracing.h:14 */
};

/* Device buffers reachable through the test seam (mrt_read_buffer / mrt_write_buffer). */
enum {
    MRT_BUF_RAYS = 0,            /* mrt_ray[W*H]                 Renderer.mm:651 */
    MRT_BUF_LIGHT_RAYS = 1,      /* mrt_light_sampling_ray[W*H]  Renderer.mm:652 */
    MRT_BUF_INTERSECTIONS = 2,   /* mrt_intersection[W*H]        Renderer.mm:653 */
    MRT_BUF_IMAGE = 3            /* float4[W*H], row 0 = bottom  Renderer.mm:645-648 */
};

typedef struct mrt_ctx mrt_ctx;
typedef struct mrt_scene mrt_scene;

/* Compile-time macros of the reference (Raytracing.h:11-25) turned into run-time configuration.
 * mrt_config_default() fills in the reference's values. */
typedef struct mrt_config {
    uint32_t struct_size;       /* = sizeof(mrt_config); checked */
    int32_t  device;            /* CUDA device ordinal */
    uint32_t max_path_length;   /* MAX_PATH_LENGTH, Raytracing.h:23 (1..64) */
    uint32_t accumulate_mode;   /* MRT_ACCUM_*; default MEAN (ACCUMULATE_IMAGE true) */
    uint32_t animate_noise;     /* ANIMATE_NOISE, Raytracing.h:20: frame f refreshes slot f%3 */
    uint32_t pipeline;          /* MRT_PIPELINE_* */
    uint32_t bvh_max_leaf;      /* SAH builder: max triangles per leaf (default 4) */
    uint32_t bvh_threads;       /* SAH builder host threads (0 = hardware concurrency) */
    uint32_t count_rays;        /* 1: maintain live-ray counters (mrt_stats) */
    uint32_t reserved[7];
} mrt_config;

typedef struct mrt_stats {
    uint64_t primary_rays;      /* live path segments traversed (maxDistance >= 0 at intersect time) */
    uint64_t shadow_rays;       /* live light-sampling rays traversed */
    uint64_t kernel_launches;   /* kernels of this library launched since mrt_create */
    uint64_t frames_rendered;
    uint64_t nodes_visited;     /* filled by mrt_profile_traversal and by renders under MRT_PROFILE_VISITS */
    uint64_t triangles_tested;  /* likewise */
    float    last_render_ms;    /* device time of the last mrt_render_frames (CUDA events) */
    uint32_t bvh_nodes;
    uint32_t bvh_depth;
    float    bvh_sah_cost;
    float    bvh_build_ms;      /* host wall time of the last SAH build */
    float    last_kernel_ms;    /* device time of the kernels alone inside the last mrt_render_frames */
    uint32_t bvh8_nodes;        /* 8-wide quantised tree the production trace kernel traverses (80-byte nodes) */
    uint32_t bvh8_depth;
    float    profile_trace_ms;  /* mrt_set_profile(1): summed CUDA-event time of the k_wf_trace launches of the last mrt_render_frames */
    uint32_t profile_trace_launches;   /* ... and how many of them found work (queue not empty) */
    float    bvh8_build_ms;     /* host wall time of the wide-tree build (all SAH-bin candidates + the sampled selection) */
    uint32_t waves_traced;      /* trace launches of the last wavefront batch that found rays */
    uint32_t tail_waves;        /* profile mode: waves of the last mrt_render_frames with fewer extension rays than half the path slots */
    uint32_t reserved;
} mrt_stats;

/* Read-only view of a host scene container (arrays in the ABI layouts of mrt_abi.h). */
typedef struct mrt_scene_view {
    const mrt_vertex* vertices;               uint32_t vertex_count;
    const uint32_t* indices;                  uint32_t triangle_count;   /* 3 per triangle */
    const mrt_triangle_reference* references;                          /* triangle_count */
    const mrt_material* materials;            uint32_t material_count;
    const mrt_light_triangle* lights;         uint32_t light_count;      /* array has light_count+1 */
} mrt_scene_view;

/* ---- errors / version ------------------------------------------------------------------- */
const char* mrt_last_error(void);
int  mrt_version(void);
void mrt_config_default(mrt_config* cfg);

/* ---- lifecycle -------------------------------------------------------------------------- */
/* -[Renderer initWithMetalKitView:] + loadMetalWithView: (Renderer.mm:72-154): device, queue,
 * per-frame SharedData and the three noise slots (zero-filled until mrt_set_noise). */
int mrt_create(const mrt_config* cfg, mrt_ctx** out);
int mrt_destroy(mrt_ctx* ctx);

/* Buffer upload + acceleration structure of -initRaytracing (Renderer.mm:450-469): vertex
 * (stride 24) / uint32 index / TriangleReference / Material / LightTriangle buffers, then a SAH BVH
 * built on the host replaces MPSTriangleAccelerationStructure -rebuild.  `lights` holds
 * light_count+1 entries (sentinel last, Renderer.mm:444-448). */
int mrt_set_scene(mrt_ctx* ctx,
                  const mrt_vertex* vertices, uint32_t vertex_count,
                  const uint32_t* indices, uint32_t triangle_count,
                  const mrt_triangle_reference* references,
                  const mrt_material* materials, uint32_t material_count,
                  const mrt_light_triangle* lights, uint32_t light_count);
int mrt_set_scene_from(mrt_ctx* ctx, const mrt_scene* scene);
/* The same in two halves, for several contexts (GPUs) in one process: the host half — validation, both SAH trees, the
 * records repacked for 128-bit loads — runs once, the device half uploads to each ctx.  A prepared scene is read-only
 * after creation: any number of threads may upload from it at the same time. */
typedef struct mrt_prepared_scene mrt_prepared_scene;
int  mrt_scene_prepare(const mrt_vertex* vertices, uint32_t vertex_count,
                       const uint32_t* indices, uint32_t triangle_count,
                       const mrt_triangle_reference* references,
                       const mrt_material* materials, uint32_t material_count,
                       const mrt_light_triangle* lights, uint32_t light_count,
                       uint32_t bvh_max_leaf, uint32_t bvh_threads, mrt_prepared_scene** out);
int  mrt_scene_prepare_from(const mrt_scene* scene, uint32_t bvh_threads, mrt_prepared_scene** out);
int  mrt_set_prepared_scene(mrt_ctx* ctx, const mrt_prepared_scene* prepared);
void mrt_prepared_scene_destroy(mrt_prepared_scene* prepared);

/* -mtkView:drawableSizeWillChange: (Renderer.mm:640-657): (re)allocate Ray / LightSamplingRay /
 * Intersection / image for width x height rays, zero-fill, restart accumulation.  width/height are
 * the DISPATCH size (the reference's CONTENT_SCALE x drawable size). */
int mrt_resize(mrt_ctx* ctx, uint32_t width, uint32_t height);
int mrt_get_size(mrt_ctx* ctx, uint32_t* width, uint32_t* height);   /* 0 x 0 before the first mrt_resize */

/* Noise slot upload (Renderer.mm:115-129 initial fill, 492-496 per-frame refresh).
 * slot in [0,3); tile = 64*64 float4 in [0,1). */
int mrt_set_noise(mrt_ctx* ctx, uint32_t slot, const float* tile);

/* Multi-GPU row sharding: this ctx renders only rows y with y % row_stride == first_row.
 * (0,1) = whole image.  No reference counterpart (single device). */
int mrt_set_row_shard(mrt_ctx* ctx, uint32_t first_row, uint32_t row_stride);
int mrt_set_accumulate_mode(mrt_ctx* ctx, uint32_t mode);
int mrt_set_pipeline(mrt_ctx* ctx, uint32_t pipeline);
/* How the wavefront pipeline issues its waves (same kernels, same bits).  MRT_SCHEDULE_GRAPH (default): one CUDA graph per
 * batch whose WHILE node repeats {trace, shade, trace, shade} until the trace kernel finds the batch finished — no empty
 * launches, one cudaGraphLaunch from the host.  MRT_SCHEDULE_HOST_LOOP: the host launches the waves itself, at most 16
 * ahead of the device, and stops when the trace kernel has flagged the end in a mapped host word; this is also what runs
 * under MRT_PROFILE_TRACE_TIME, and what a kernel profiler needs (ncu cannot see kernel nodes of conditional graphs). */
enum { MRT_SCHEDULE_GRAPH = 0, MRT_SCHEDULE_HOST_LOOP = 1 };
int mrt_set_wave_schedule(mrt_ctx* ctx, uint32_t schedule);

/* ---- framebuffer reduction over NVLink peer memory (the reference is single-device) -----------------
 * Sample sharding leaves one SUM image per GPU (MRT_ACCUM_SUM); the job's image is scale * (their sum).  Instead of
 * copy + ncclAllReduce + scale, every rank enqueues ONE kernel on its ctx stream that pushes slice g of its image into
 * rank g's exchange block over NVLink, waits (flags in peer memory) until all ranks have pushed, sums its own slice in
 * rank order (deterministic: every rank, every run, the same bits), scales, stores the slice into every rank's result
 * image, and ends only when every peer has done the same (csrc/peer.cu).  No NCCL, no host synchronisation; peers only
 * ever touch a rank's exchange block, never its SUM image.
 *   mrt_peer_export    allocates this ctx's exchange block (flags + result + staging, 2 x the image) and fills the
 *                      handle the other ranks need.  sharing = MRT_PEER_SHARE_PROCESS: all ranks live in this process
 *                      (one thread per GPU): the block is mapped on every GPU through the driver's virtual memory
 *                      management, no cudaDeviceEnablePeerAccess.  MRT_PEER_SHARE_IPC: ranks are processes (torchrun):
 *                      a cudaMalloc block with a CUDA IPC handle (also usable inside one process).
 *   mrt_peer_connect   handles[world] in rank order, handles[rank] being this ctx's own.  Every rank must have
 *                      connected before any rank reduces (barrier in the caller).  mrt_resize drops block and connection.
 *   mrt_peer_reduce    asynchronous; all ranks call it once per step, in the same order.
 *   mrt_peer_result    this rank's copy of the reduced image (same bits on every rank).
 *   mrt_peer_status    synchronises the ctx stream; MRT_E_STATE if a wait gave up (~2 s without the peer showing up).
 *   mrt_peer_disconnect closes this rank's mappings of its peers' blocks (its own block lives until mrt_resize /
 *                      mrt_destroy).  Between processes: every rank disconnects, THEN a barrier, before any rank resizes
 *                      or destroys its context — freeing exported memory under an importer is undefined in CUDA. */
#define MRT_PEER_MAX_RANKS 16
enum { MRT_PEER_SHARE_IPC = 0, MRT_PEER_SHARE_PROCESS = 1 };
typedef struct mrt_peer_handle {
    unsigned char block_ipc[64];                  /* cudaIpcMemHandle_t of the exchange block (ipc_valid) */
    uint64_t block_ptr;                           /* the block as a device pointer, valid inside the exporting process */
    int64_t pid;
    int32_t device;
    uint32_t width, height, ipc_valid, mapped_everywhere, reserved;
} mrt_peer_handle;
int mrt_peer_export(mrt_ctx* ctx, uint32_t sharing, mrt_peer_handle* out);
int mrt_peer_connect(mrt_ctx* ctx, uint32_t rank, uint32_t world, const mrt_peer_handle* handles);
int mrt_peer_reduce(mrt_ctx* ctx, float scale);
int mrt_peer_result(mrt_ctx* ctx, void** device_ptr, size_t* bytes);
int mrt_peer_status(mrt_ctx* ctx);
int mrt_peer_read_result(mrt_ctx* ctx, float* rgba);   /* mrt_peer_status, then width * height * 4 floats to the host */
int mrt_peer_disconnect(mrt_ctx* ctx);

/* ---- optics extension: SURVEY.md 8(f4), BEYOND THE REFERENCE ------------------------------------------------------------
 * Nothing here has a counterpart in the reference, so nothing here can be pinned against it: the checker is the same
 * algorithm restated in oracle/ (bit-exact against the CUDA path) plus closed forms (Snell's law through a slab, total
 * internal reflection, the alias table's marginals).  flags = 0 (the default) is the reference's behaviour bit for bit
 * and launches the same kernels as before; any other value launches separate instantiations of the shade kernels.
 *   MRT_EXT_REFRACTION   smooth dielectrics refract by Snell's law instead of continuing straight through
 *                        (KernelHelpers.h:151-166 keeps wO = wI): side from the sign of n.wI, total internal reflection,
 *                        both lobes weigh 1, radiance scaled by (etaI/etaT)^2 on transmission, the continued path and
 *                        its light-sampling ray start on the side the path leaves on.  No nested media.
 *   MRT_EXT_DISPERSION   (needs REFRACTION) the index of refraction of channel c in {0,1,2} is material.ior +
 *                        dispersion * (c - 1); a path entering a medium keeps ONE live channel (uniform choice, weight =
 *                        number of live channels).  The 3 channels are Spectrum.h:3's SPECTRUM_SAMPLES; more samples
 *                        than 3 would change every ABI struct that embeds a Spectrum and is NOT built.
 *   MRT_EXT_CONTINUOUS_SPECTRUM (needs DISPERSION) the index is taken uniformly inside the kept channel's band,
 *                        material.ior + dispersion * (c - 1 + u - 0.5), u uniform: refraction sampled at a continuum of
 *                        wavelengths (no three separate ghost images) while radiometry stays 3 samples wide.
 *   MRT_EXT_ALIAS_LIGHTS light triangles are selected through a Walker/Vose alias table of the same pdfs (2 loads)
 *                        instead of the CDF search of KernelHelpers.h:49-54; same expectation, different samples.
 * Wavefront and staged pipelines; MRT_E_STATE under MRT_PIPELINE_MEGAKERNEL.  The setting belongs to the ctx, survives
 * mrt_set_scene / mrt_resize and is NOT stored by mrt_checkpoint_save: set it again after mrt_checkpoint_load. */
enum { MRT_EXT_REFRACTION = 1, MRT_EXT_DISPERSION = 2, MRT_EXT_ALIAS_LIGHTS = 4, MRT_EXT_CONTINUOUS_SPECTRUM = 8 };
int mrt_set_extensions(mrt_ctx* ctx, uint32_t flags, float dispersion);
/* The alias table mrt_set_scene builds from the light pdfs (host code, no GPU): slot i keeps itself when the fraction
 * is below prob[i], else yields alias[i]. */
int mrt_build_alias_table(const float* pdf, uint32_t count, float* prob, uint32_t* alias);
/* Profiling flags (OR them).  MRT_PROFILE_TRACE_TIME: bracket every trace-kernel launch of mrt_render_frames (wavefront
 * pipeline) with a CUDA event pair on the ctx stream, so that mrt_get_stats can report the dominant kernel's own device
 * time (costs two event records per wave).  MRT_PROFILE_VISITS: run the counting variant of the trace kernel, which adds
 * the node tests and triangle tests of every ray into mrt_stats.nodes_visited / triangles_tested (needs count_rays;
 * slower — for the BVH term of the roofline, never inside a timed region).  Same bits in the image either way. */
#define MRT_PROFILE_TRACE_TIME 1u
#define MRT_PROFILE_VISITS     2u
int mrt_set_profile(mrt_ctx* ctx, uint32_t flags);

/* ---- the six dispatch kinds, one entry point each (asynchronous on the ctx stream) ---------- */
/* rayGenerator, Shaders.metal:75-103; bindings Renderer.mm:508-510 (noise slot frame%3). */
int mrt_raygen(mrt_ctx* ctx, uint32_t frame_index);
/* MPSRayIntersector nearest hit, Renderer.mm:519-523 (which = MRT_BUF_RAYS, stride 80) and
 * 545-553 (which = MRT_BUF_LIGHT_RAYS, stride 48); writes the single shared Intersection buffer.
 * Semantics of Renderer.mm:464-469: no culling, no masks, rays with maxDistance < 0 are skipped and
 * report distance < 0. */
int mrt_intersect(mrt_ctx* ctx, uint32_t which);
/* intersectionHandler, Shaders.metal:105-212; bindings Renderer.mm:529-539
 * (noise slot (frame+bounce_iteration)%3). */
int mrt_shade(mrt_ctx* ctx, uint32_t frame_index, uint32_t bounce_iteration);
/* lightSamplingHandler, Shaders.metal:214-231; bindings Renderer.mm:559-565. */
int mrt_shadow_resolve(mrt_ctx* ctx);
/* accumulateImage, Shaders.metal:233-249; bindings Renderer.mm:577-579. */
int mrt_accumulate(mrt_ctx* ctx, uint32_t frame_index);

/* -performRaytracing: (Renderer.mm:500-585) as the reference sequences it: raygen, then
 * max_path_length x {intersect, shade, intersect, resolve}, then accumulate — 2 + 4*L launches
 * through the full-size stage buffers.  Does not touch the noise slots. */
int mrt_render_frame_staged(mrt_ctx* ctx, uint32_t frame_index);

/* The production path: frames first_frame, first_frame+frame_stride, ... (n_frames of them)
 * rendered by the fused sm_100a pipeline, bit-identical to the staged sequence.
 * noise_tiles: host array of (n_frames-1)*frame_stride+1 tiles, tile k belongs to frame
 * first_frame+k and refreshes slot (first_frame+k)%3 before that frame is rendered, exactly as
 * -updateSharedData does (Renderer.mm:485-497).  NULL = no refresh (ANIMATE_NOISE 0).  On return
 * the three slots hold what the reference's would after the last frame.  noise_tiles may be a host pointer
 * (pageable: borrowed until the call returns; page-locked: must stay unchanged until mrt_sync) or a device
 * pointer on the ctx's device (inputs already resident in HBM).  Asynchronous. */
int mrt_render_frames(mrt_ctx* ctx, uint32_t first_frame, uint32_t n_frames, uint32_t frame_stride,
                      const float* noise_tiles);
int mrt_sync(mrt_ctx* ctx);
/* Device-side timing on the ctx stream (CUDA events): record marker `slot` (0..7) now / milliseconds between two
 * recorded markers (synchronises on the later one).  The reference only has the wall-clock frame time of
 * Renderer.mm:474-476, 631-637. */
int mrt_event_record(mrt_ctx* ctx, uint32_t slot);
int mrt_event_elapsed_ms(mrt_ctx* ctx, uint32_t slot_begin, uint32_t slot_end, float* ms);

/* ---- results / test seams ------------------------------------------------------------------- */
/* Image read-back (the reference never reads back; -saveCurrentImage is a stub, Renderer.mm:659-662).
 * rgba: width*height*4 floats, row 0 = bottom row. */
int mrt_read_image(mrt_ctx* ctx, float* rgba);
int mrt_scale_image(mrt_ctx* ctx, float scale);          /* rgb *= scale, alpha = 1 */
int mrt_clear_image(mrt_ctx* ctx);
int mrt_image_device_ptr(mrt_ctx* ctx, void** device_ptr, size_t* bytes); /* for NCCL reduce */

/* ---- multi-GPU plumbing (no reference counterpart: one MTLCommandQueue on one device, Renderer.mm:153) -------------
 * The library itself never talks to NCCL; the caller (host/mrt_render.c, bench.py) enqueues its collective on the
 * ctx stream between these calls, so a step needs no host synchronisation.
 *   sample sharding: MRT_ACCUM_SUM renders -> ncclAllReduce(image -> result) -> mrt_scale_buffer(result, 1/frames)
 *   row sharding:    mrt_set_row_shard     -> mrt_pack_rows(send) -> ncclAllGather(send -> recv) -> mrt_unpack_rows(recv)
 * mrt_stream_handle returns the ctx's cudaStream_t.  mrt_scale_buffer: rgb *= scale, alpha = 1 on any device buffer of
 * `pixels` float4.  mrt_shard_layout: rows_max = ceil(H / shards) and the bytes one shard sends.  mrt_pack_rows copies
 * the rows this ctx owns (mrt_set_row_shard) to a contiguous device buffer of rows_max * W float4; mrt_unpack_rows
 * scatters `shards` such blocks (shard g at offset g * bytes_per_shard) back to rows g, g + shards, ... of the image —
 * bit-identical to the single-device image, and the owned rows keep accumulating across steps. */
int mrt_stream_handle(mrt_ctx* ctx, void** cuda_stream);
int mrt_scale_buffer(mrt_ctx* ctx, void* device_rgba, size_t pixels, float scale);
int mrt_shard_layout(mrt_ctx* ctx, uint32_t shards, uint32_t* rows_max, size_t* bytes_per_shard);
int mrt_pack_rows(mrt_ctx* ctx, void* device_dst);
int mrt_unpack_rows(mrt_ctx* ctx, const void* device_src, uint32_t shards);
int mrt_read_buffer(mrt_ctx* ctx, uint32_t which, void* dst, size_t bytes);
int mrt_write_buffer(mrt_ctx* ctx, uint32_t which, const void* src, size_t bytes);
int mrt_get_stats(mrt_ctx* ctx, mrt_stats* out);
int mrt_reset_stats(mrt_ctx* ctx);

/* ---- presentation pass and checkpoint: the callers' side of the path ---- */
/* blitFragment (Shaders.metal:33-69) on the headless image: tone mapping 1-exp(-c) (ENABLE_TONE_MAPPING,
 * Raytracing.h:11), manual sRGB encode (MANUAL_SRGB, Raytracing.h:12; toSRGB Raytracing.h:130-135) and the comparison
 * modes of Raytracing.h:27-33 against a reference image (nearest texel, any size; Renderer.mm:160-250 loads it).
 * The macros of the reference are run-time parameters here; mrt_display_params_default gives the reference's values. */
typedef struct mrt_display_params {
    uint32_t tone_mapping;      /* 0 / 1 */
    uint32_t manual_srgb;       /* 0 / 1 */
    uint32_t comparison_mode;   /* MRT_COMPARE_* (mrt_abi.h) */
    float comparison_scale;     /* COMPARISON_SCALE */
} mrt_display_params;
void mrt_display_params_default(mrt_display_params* p);
/* out_rgba: width*height*4 floats (host), row 0 = bottom.  reference_rgba: ref_width*ref_height*4 floats (host or
 * device), row 0 = bottom, or NULL (comparison disabled whatever the mode). */
int mrt_display(mrt_ctx* ctx, const mrt_display_params* params, const float* reference_rgba, uint32_t ref_width,
                uint32_t ref_height, float* out_rgba);

/* Accumulation checkpoint: image, size, frames accumulated, accumulate mode and the three noise slots — what a long
 * progressive render needs to continue bit-exactly in another process (the reference only has the stub
 * -saveCurrentImage, Renderer.mm:626-629, 659-662).  load resizes the context to the stored size. */
int mrt_checkpoint_save(mrt_ctx* ctx, const char* path);
int mrt_checkpoint_load(mrt_ctx* ctx, const char* path, uint32_t* frames_accumulated);

/* Stand-alone replacement of one MPSRayIntersector encode (Renderer.mm:464-469, 519-523): `rays`
 * is a HOST array of `count` records of `stride_bytes` each whose first 32 bytes are
 * origin[3], minDistance, direction[3], maxDistance; `out` receives count intersections. */
int mrt_intersect_host(mrt_ctx* ctx, const void* rays, size_t stride_bytes, uint32_t count,
                       mrt_intersection* out);
/* mrt_intersect_host WITHOUT any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same ray convention and the same results bit for bit; `rays` and `out` may be host or
 * device pointers and are borrowed until the call returns (it synchronises the ctx stream).  The scratch
 * mrt_intersect_host allocates per call (staged rays, hit records) and the traversal's work counter live in ONE static
 * device array of MRT_NOALLOC_SCRATCH_BYTES (6 MiB + 256 B) that every GPU loading the library holds until the process
 * exits; its device address is resolved once per device on the first call there (cudaGetSymbolAddress).  Limits:
 * count <= MRT_NOALLOC_MAX_RAYS and stride_bytes <= MRT_NOALLOC_MAX_STRIDE, else MRT_E_INVALID with a message naming the
 * limit.  The call issues cudaMemsetAsync / cudaMemcpyAsync / one kernel launch / cudaStreamSynchronize — no cudaMalloc,
 * cudaFree or allocating library routine.  ALL calls on one device share the array and it keeps the last call's
 * contents (the call clears the work counter itself): calls that overlap on the GPU — two contexts or two threads on
 * the same device — RACE; serialise them. */
#define MRT_NOALLOC_MAX_RAYS   65536u   /* rays per call */
#define MRT_NOALLOC_MAX_STRIDE 80u      /* bytes per ray record (= sizeof(mrt_ray), the stride of Renderer.mm:519) */
#define MRT_NOALLOC_SCRATCH_BYTES (256u + MRT_NOALLOC_MAX_RAYS * MRT_NOALLOC_MAX_STRIDE + MRT_NOALLOC_MAX_RAYS * 16u)
int mrt_intersect_host_noalloc(mrt_ctx* ctx, const void* rays, size_t stride_bytes, uint32_t count,
                               mrt_intersection* out);
/* Same rays through an instrumented traversal; adds node/triangle visit counts to mrt_stats. */
int mrt_profile_traversal(mrt_ctx* ctx, const void* rays, size_t stride_bytes, uint32_t count);

/* ---- caller-side helpers (host only; usable without a GPU) ---------------------------------- */
/* Scene -> ABI buffers, the job of -initRaytracing before upload (Renderer.mm:257-448):
 * materials, triangle references, light list with pdf/cdf and sentinel. */
int  mrt_scene_create_procedural(const char* name, uint64_t seed, mrt_scene** out);
int  mrt_scene_load_obj(const char* obj_path, mrt_scene** out);
int  mrt_scene_get(const mrt_scene* scene, mrt_scene_view* view);
void mrt_scene_destroy(mrt_scene* scene);
/* Host-only self-test of the acceleration structure (no GPU needed): builds the 8-wide BVH the kernels traverse,
 * sends n_rays seeded random rays (some axis-parallel, some starting on box planes) through a host copy of the
 * traversal state machine — same node decode as the CUDA kernel — and compares every (distance, triangle) with brute
 * force over all triangles.  Stands where MPSTriangleAccelerationStructure would simply be trusted (Renderer.mm:456-462). */
typedef struct mrt_bvh_report {
    uint32_t nodes, depth, triangles, mismatches;
    float    sah_cost, build_ms;
    float    nodes_per_ray, triangles_per_ray;
} mrt_bvh_report;
int  mrt_bvh_selftest(const mrt_scene* scene, uint32_t n_rays, uint64_t seed, mrt_bvh_report* out);
/* Noise tile for (seed, frame): 64*64*4 floats, (mt19937_64(seed ^ frame)() >> 40) * 2^-24.
 * Stands in for the clock-seeded fills of Renderer.mm:109-126, 486-495. */
void mrt_noise_fill(uint64_t seed, uint32_t frame, float* tile);

#ifdef __cplusplus
}
#endif
#endif /* MRT_H */
