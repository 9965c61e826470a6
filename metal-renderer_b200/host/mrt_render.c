/*
 * mrt_render.c — headless plain-C driver over the C-ABI (include/mrt.h), one or several GPUs.
 *
 * Replaces the windowed frame loop of the reference (-drawInMTKView:, renderer/Renderer.mm:587-638, driving
 * -performRaytracing: :500-585, and the NSApplication shells under macos/ and ios/): build or load a scene, render N
 * frames, write the accumulated image as a PFM file, print one JSON line with the reference's own statistic
 * ("Mrays/s" = pixel-samples per second, Renderer.mm:631-637) plus live ray segments per second.
 *
 * Multi-GPU (--devices a,b,...; built with -DMRT_WITH_NCCL): one process, one host thread and one mrt_ctx per GPU,
 * the scene prepared ONCE on the host (mrt_scene_prepare_from) and uploaded to every GPU, ncclCommInitAll, and ONE
 * collective per render on each ctx's own stream (mrt_stream_handle):
 *   --shard samples (default): GPU g renders frames g, g+N, ... into a sum image (MRT_ACCUM_SUM);
 *                    ncclAllReduce(sum image -> result) over NVLink, then mrt_scale_buffer(result, 1/frames).
 *   --shard rows:    GPU g renders rows y % N == g of every frame with the reference's running mean;
 *                    mrt_pack_rows -> ncclAllGather -> mrt_unpack_rows.  Bit-identical to one GPU.
 *   --reduce peer:   (samples) no NCCL at all: mrt_peer_reduce, one kernel per GPU that sums the images over NVLink peer
 *                    memory in rank order, scales and publishes the result to every GPU (csrc/peer.cu).
 * The library itself never links NCCL.
 *
 * usage: mrt_render [--scene NAME|file.obj] [--size WxH] [--frames N] [--depth L] [--seed S] [--device D]
 *                   [--devices a,b,...] [--shard samples|rows] [--reduce nccl|peer]
 *                   [--pipeline auto|wavefront|fused|staged] [--out image.pfm]
 *                   [--tone] [--srgb]            presentation pass on the written image (Shaders.metal:43-51; one GPU)
 *                   [--resume file.ckpt]         continue an earlier render (frames then count from the stored index; one GPU)
 *                   [--checkpoint file.ckpt]     save image + frame index + noise slots after rendering (one GPU)
 *                   [--optics refraction|dispersion|spectrum|alias|all] [--spread D]
 *                                                optics extension, NOT in the reference (mrt_set_extensions): Snell refraction,
 *                                                per-channel index material.ior + D * (c - 1) (default D 0.02), alias-table lights
 */
#define _POSIX_C_SOURCE 200809L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "mrt.h"
#ifdef MRT_WITH_NCCL
#include <pthread.h>
#include <cuda_runtime_api.h>
#include <nccl.h>
#endif

#define CHECK(call) do { int rc_ = (call); if (rc_ != MRT_OK) { \
    fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, mrt_last_error()); return 1; } } while (0)

static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static int write_pfm(const char* path, const float* rgba, unsigned w, unsigned h) {
    FILE* f = fopen(path, "wb");
    if (!f) return -1;
    fprintf(f, "PF\n%u %u\n-1.0\n", w, h);            /* little endian; PFM rows run bottom to top, like the image */
    for (size_t i = 0; i < (size_t)w * h; ++i) fwrite(rgba + 4 * i, sizeof(float), 3, f);
    fclose(f);
    return 0;
}

typedef struct options {
    const char* scene_name; const char* out; const char* resume; const char* checkpoint;
    unsigned tone, srgb, w, h, frames, depth, pipeline, rows, optics, peer;
    float spread;
    int devices[16]; int n_devices;
    unsigned long long seed;
} options;

#ifdef MRT_WITH_NCCL
/* ---- one thread per GPU ------------------------------------------------------------------------------------------ */
typedef struct worker {
    const options* opt; const mrt_prepared_scene* prepared; const float* tiles; const float* init;
    int rank, world; ncclComm_t comm; pthread_barrier_t* barrier;
    mrt_peer_handle* handles;      /* --reduce peer: one per rank, shared by all workers */
    float* image;                  /* rank 0: host copy of the job's image */
    mrt_stats stats; float device_ms, collective_ms; double setup_ms, teardown_ms; int rc, passed, barriers; char error[512];
} worker;

#define WCHECK(call) do { int rc_ = (call); if (rc_ != MRT_OK) { snprintf(w->error, sizeof w->error, "%s failed (%d): %s", #call, rc_, mrt_last_error()); \
    w->rc = 1; goto done; } } while (0)
#define WCUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { snprintf(w->error, sizeof w->error, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    w->rc = 1; goto done; } } while (0)
#define WNCCL(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) { snprintf(w->error, sizeof w->error, "%s failed: %s", #call, ncclGetErrorString(r_)); \
    w->rc = 1; goto done; } } while (0)

static void* worker_main(void* arg) {
    worker* w = (worker*)arg;
    const options* o = w->opt;
    const unsigned N = (unsigned)w->world, g = (unsigned)w->rank;
    mrt_ctx* ctx = NULL;
    void* result = NULL; void* send = NULL; void* recv = NULL; void* stream = NULL; void* peer_result = NULL;
    void* image_dev = NULL; size_t image_bytes = 0;
    mrt_config cfg;
    const double t_start = now_s();
    mrt_config_default(&cfg);
    cfg.device = o->devices[g];
    cfg.max_path_length = o->depth;
    cfg.pipeline = o->pipeline;
    cfg.accumulate_mode = o->rows ? MRT_ACCUM_MEAN : MRT_ACCUM_SUM;
    WCHECK(mrt_create(&cfg, &ctx));
    WCHECK(mrt_set_prepared_scene(ctx, w->prepared));
    if (o->optics) WCHECK(mrt_set_extensions(ctx, o->optics, o->spread));
    WCHECK(mrt_resize(ctx, o->w, o->h));
    WCHECK(mrt_stream_handle(ctx, &stream));
    WCHECK(mrt_image_device_ptr(ctx, &image_dev, &image_bytes));
    for (unsigned s = 0; s < MRT_NOISE_SLOTS; ++s) WCHECK(mrt_set_noise(ctx, s, w->init));
    const size_t pixels = (size_t)o->w * o->h;
    if (o->rows) {
        uint32_t rows_max = 0; size_t shard_bytes = 0;
        WCHECK(mrt_set_row_shard(ctx, g, N));
        WCHECK(mrt_shard_layout(ctx, N, &rows_max, &shard_bytes));
        WCUDA(cudaMalloc(&send, shard_bytes));
        WCUDA(cudaMalloc(&recv, shard_bytes * N));
        WNCCL(ncclAllGather(send, recv, 32, ncclFloat, w->comm, (cudaStream_t)stream));   /* untimed: NCCL sets its channels up on first use */
        WCHECK(mrt_sync(ctx));
        pthread_barrier_wait(w->barrier); w->passed++;
        w->setup_ms = (now_s() - t_start) * 1e3;
        WCHECK(mrt_event_record(ctx, 0));
        WCHECK(mrt_render_frames(ctx, 0, o->frames, 1, w->tiles));            /* every frame, owned rows only */
        WCHECK(mrt_pack_rows(ctx, send));
        WCHECK(mrt_event_record(ctx, 1));
        WNCCL(ncclAllGather(send, recv, shard_bytes / sizeof(float), ncclFloat, w->comm, (cudaStream_t)stream));
        WCHECK(mrt_event_record(ctx, 2));
        WCHECK(mrt_unpack_rows(ctx, recv, N));
        WCHECK(mrt_event_record(ctx, 3));
        if (g == 0) WCHECK(mrt_read_image(ctx, w->image)); else WCHECK(mrt_sync(ctx));
    } else {
        /* frames g, g+N, ...: the slots of the two frames before this rank's first one must hold their tiles
         * (slot f%3 is refreshed at frame f, bounce i reads slot (f+i)%3: Renderer.mm:492-496, 538) */
        const unsigned mine = o->frames > g ? (o->frames - g + N - 1) / N : 0;
        for (unsigned f = g >= 2 ? g - 2 : 0; f < g; ++f) WCHECK(mrt_set_noise(ctx, f % 3, w->tiles + (size_t)f * MRT_NOISE_TILE_FLOATS));
        if (o->peer) {                                  /* one kernel per GPU over NVLink peer memory instead of NCCL (csrc/peer.cu) */
            const double p0 = now_s();
            WCHECK(mrt_peer_export(ctx, MRT_PEER_SHARE_PROCESS, &w->handles[g]));
            const double p1 = now_s();
            pthread_barrier_wait(w->barrier); w->passed++;                 /* every handle is there */
            const double p2 = now_s();
            WCHECK(mrt_peer_connect(ctx, g, N, w->handles));
            const double p3 = now_s();
            pthread_barrier_wait(w->barrier); w->passed++;                 /* nobody reduces before everybody has connected */
            WCHECK(mrt_peer_result(ctx, &peer_result, NULL));
            WCHECK(mrt_peer_reduce(ctx, 0.0f));             /* untimed, like NCCL's first call below: first touch of every peer's pages */
            WCHECK(mrt_peer_status(ctx));
            if (getenv("MRT_RENDER_VERBOSE"))
                fprintf(stderr, "rank %u: before export %.1f ms, export %.1f, wait %.1f, connect %.1f, barrier + first reduce %.1f\n", g,
                        (p0 - t_start) * 1e3, (p1 - p0) * 1e3, (p2 - p1) * 1e3, (p3 - p2) * 1e3, (now_s() - p3) * 1e3);
        } else {
            WCUDA(cudaMalloc(&result, image_bytes));
            WNCCL(ncclAllReduce(result, result, 32, ncclFloat, ncclSum, w->comm, (cudaStream_t)stream));   /* untimed: first-use set-up */
        }
        WCHECK(mrt_sync(ctx));
        pthread_barrier_wait(w->barrier); w->passed++;
        w->setup_ms = (now_s() - t_start) * 1e3;
        WCHECK(mrt_event_record(ctx, 0));
        if (mine) WCHECK(mrt_render_frames(ctx, g, mine, N, w->tiles + (size_t)g * MRT_NOISE_TILE_FLOATS));
        WCHECK(mrt_event_record(ctx, 1));
        if (o->peer) {
            WCHECK(mrt_peer_reduce(ctx, 1.0f / (float)o->frames));
            WCHECK(mrt_event_record(ctx, 2));
            WCHECK(mrt_event_record(ctx, 3));
            WCHECK(mrt_peer_status(ctx));                                  /* synchronises; a wait that gave up is an error */
        } else {
            WNCCL(ncclAllReduce(image_dev, result, pixels * 4, ncclFloat, ncclSum, w->comm, (cudaStream_t)stream));
            WCHECK(mrt_event_record(ctx, 2));
            WCHECK(mrt_scale_buffer(ctx, result, pixels, 1.0f / (float)o->frames));
            WCHECK(mrt_event_record(ctx, 3));
            WCHECK(mrt_sync(ctx));
        }
        if (g == 0) WCUDA(cudaMemcpy(w->image, o->peer ? peer_result : result, image_bytes, cudaMemcpyDeviceToHost));
    }
    WCHECK(mrt_event_elapsed_ms(ctx, 0, 3, &w->device_ms));
    WCHECK(mrt_event_elapsed_ms(ctx, 1, 2, &w->collective_ms));
    WCHECK(mrt_get_stats(ctx, &w->stats));
done:
    while (w->rc && w->passed < w->barriers) { pthread_barrier_wait(w->barrier); w->passed++; }   /* never leave the others waiting */
    {
        const double t_end = now_s();
        if (result) cudaFree(result);
        if (send) cudaFree(send);
        if (recv) cudaFree(recv);
        if (ctx) mrt_destroy(ctx);
        w->teardown_ms = (now_s() - t_end) * 1e3;
    }
    return NULL;
}

static int render_multi(const options* o, const mrt_scene* scene, const mrt_scene_view* view) {
    const int N = o->n_devices;
    mrt_prepared_scene* prepared = NULL;
    double tb = now_s();
    CHECK(mrt_scene_prepare_from(scene, 0, &prepared));            /* both trees built once, on all host cores */
    double build_ms = (now_s() - tb) * 1e3;
    const int use_peer = o->peer && !o->rows;
    ncclComm_t comms[16];
    memset(comms, 0, sizeof comms);
    mrt_peer_handle handles[16];
    if (!use_peer) {
        ncclResult_t nr = ncclCommInitAll(comms, N, o->devices);
        if (nr != ncclSuccess) { fprintf(stderr, "ncclCommInitAll failed: %s\n", ncclGetErrorString(nr)); return 1; }
    }
    const unsigned frames = o->frames ? o->frames : 1;
    float* tiles = (float*)malloc((size_t)frames * MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* init = (float*)malloc(MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* image = (float*)malloc((size_t)o->w * o->h * 4 * sizeof(float));
    if (!tiles || !init || !image) { fprintf(stderr, "out of memory\n"); return 1; }
    mrt_noise_fill(o->seed, 0xFFFFFFFFu, init);
    for (unsigned f = 0; f < frames; ++f) mrt_noise_fill(o->seed, f, tiles + (size_t)f * MRT_NOISE_TILE_FLOATS);
    pthread_barrier_t barrier;
    pthread_barrier_init(&barrier, NULL, (unsigned)N);
    worker ws[16]; pthread_t th[16];
    memset(ws, 0, sizeof ws);
    double t0 = now_s();
    for (int g = 0; g < N; ++g) {
        ws[g].opt = o; ws[g].prepared = prepared; ws[g].tiles = tiles; ws[g].init = init; ws[g].rank = g; ws[g].world = N;
        ws[g].comm = comms[g]; ws[g].barrier = &barrier; ws[g].image = image; ws[g].handles = handles; ws[g].barriers = use_peer ? 3 : 1;
        pthread_create(&th[g], NULL, worker_main, &ws[g]);
    }
    int failed = 0;
    for (int g = 0; g < N; ++g) { pthread_join(th[g], NULL); if (ws[g].rc) { fprintf(stderr, "device %d: %s\n", o->devices[g], ws[g].error); failed = 1; } }
    double t1 = now_s();
    for (int g = 0; g < N && !use_peer; ++g) ncclCommDestroy(comms[g]);
    pthread_barrier_destroy(&barrier);
    if (failed) return 1;
    float ms = 0, coll = 0; unsigned long long primary = 0, shadow = 0, launches = 0;
    double setup = 0, teardown = 0;
    for (int g = 0; g < N; ++g) {
        if (ws[g].setup_ms > setup) setup = ws[g].setup_ms;
        if (ws[g].teardown_ms > teardown) teardown = ws[g].teardown_ms;
        if (ws[g].device_ms > ms) ms = ws[g].device_ms;
        if (ws[g].collective_ms > coll) coll = ws[g].collective_ms;
        primary += ws[g].stats.primary_rays; shadow += ws[g].stats.shadow_rays; launches += ws[g].stats.kernel_launches;
    }
    double samples = (double)o->w * o->h * o->frames;
    printf("{\"scene\": \"%s\", \"triangles\": %u, \"lights\": %u, \"width\": %u, \"height\": %u, \"frames\": %u, \"depth\": %u, "
           "\"gpus\": %d, \"shard\": \"%s\", \"reduce\": \"%s\", \"device_ms\": %.3f, \"collective_ms\": %.3f, \"wall_ms\": %.3f, \"msamples_per_s\": %.2f, "
           "\"mrays_per_s\": %.2f, \"primary_rays\": %llu, \"shadow_rays\": %llu, \"kernel_launches\": %llu, \"scene_prepare_ms\": %.1f, \"rank_setup_ms\": %.1f, \"rank_teardown_ms\": %.1f}\n",
           o->scene_name, view->triangle_count, view->light_count, o->w, o->h, o->frames, o->depth, N, o->rows ? "rows" : "samples", o->rows ? "nccl all-gather" : (use_peer ? "peer kernel" : "nccl all-reduce"),
           ms, coll, (t1 - t0) * 1e3, samples / (ms * 1e3), (double)(primary + shadow) / (ms * 1e3), primary, shadow, launches, build_ms, setup, teardown);
    if (o->out && write_pfm(o->out, image, o->w, o->h)) { fprintf(stderr, "cannot write %s\n", o->out); return 1; }
    free(tiles); free(init); free(image);
    mrt_prepared_scene_destroy(prepared);
    return 0;
}
#endif /* MRT_WITH_NCCL */

int main(int argc, char** argv) {
    options o;
    memset(&o, 0, sizeof o);
    o.scene_name = "cornell"; o.w = 256; o.h = 256; o.frames = 16; o.depth = MRT_MAX_PATH_LENGTH; o.pipeline = MRT_PIPELINE_AUTO;
    o.seed = 1; o.n_devices = 1; o.devices[0] = 0; o.spread = 0.02f;
    unsigned first = 0;
    for (int i = 1; i < argc; ++i) {
        const char* a = argv[i];
        const char* v = i + 1 < argc ? argv[i + 1] : "";
        if (!strcmp(a, "--scene")) { o.scene_name = v; ++i; }
        else if (!strcmp(a, "--size")) { if (sscanf(v, "%ux%u", &o.w, &o.h) != 2) { fprintf(stderr, "bad --size\n"); return 2; } ++i; }
        else if (!strcmp(a, "--frames")) { o.frames = (unsigned)atoi(v); ++i; }
        else if (!strcmp(a, "--depth")) { o.depth = (unsigned)atoi(v); ++i; }
        else if (!strcmp(a, "--seed")) { o.seed = strtoull(v, NULL, 10); ++i; }
        else if (!strcmp(a, "--device")) { o.devices[0] = atoi(v); o.n_devices = 1; ++i; }
        else if (!strcmp(a, "--devices")) {
            o.n_devices = 0;
            for (const char* p = v; *p && o.n_devices < 16; ) { o.devices[o.n_devices++] = (int)strtol(p, (char**)&p, 10); if (*p == ',') ++p; else break; }
            if (o.n_devices < 1) { fprintf(stderr, "bad --devices\n"); return 2; }
            ++i;
        }
        else if (!strcmp(a, "--shard")) { o.rows = !strcmp(v, "rows"); if (!o.rows && strcmp(v, "samples")) { fprintf(stderr, "bad --shard\n"); return 2; } ++i; }
        else if (!strcmp(a, "--reduce")) { o.peer = !strcmp(v, "peer"); if (!o.peer && strcmp(v, "nccl")) { fprintf(stderr, "bad --reduce\n"); return 2; } ++i; }
        else if (!strcmp(a, "--pipeline")) {
            if (!strcmp(v, "staged")) o.pipeline = MRT_PIPELINE_STAGED;
            else if (!strcmp(v, "fused")) o.pipeline = MRT_PIPELINE_MEGAKERNEL;
            else if (!strcmp(v, "wavefront")) o.pipeline = MRT_PIPELINE_WAVEFRONT;
            else if (!strcmp(v, "auto")) o.pipeline = MRT_PIPELINE_AUTO;
            else { fprintf(stderr, "bad --pipeline\n"); return 2; }
            ++i;
        }
        else if (!strcmp(a, "--out")) { o.out = v; ++i; }
        else if (!strcmp(a, "--resume")) { o.resume = v; ++i; }
        else if (!strcmp(a, "--checkpoint")) { o.checkpoint = v; ++i; }
        else if (!strcmp(a, "--optics")) {
            if (!strcmp(v, "refraction")) o.optics = MRT_EXT_REFRACTION;
            else if (!strcmp(v, "dispersion")) o.optics = MRT_EXT_REFRACTION | MRT_EXT_DISPERSION;
            else if (!strcmp(v, "alias")) o.optics = MRT_EXT_ALIAS_LIGHTS;
            else if (!strcmp(v, "spectrum")) o.optics = MRT_EXT_REFRACTION | MRT_EXT_DISPERSION | MRT_EXT_CONTINUOUS_SPECTRUM;
            else if (!strcmp(v, "all")) o.optics = MRT_EXT_REFRACTION | MRT_EXT_DISPERSION | MRT_EXT_ALIAS_LIGHTS | MRT_EXT_CONTINUOUS_SPECTRUM;
            else { fprintf(stderr, "bad --optics\n"); return 2; }
            ++i;
        }
        else if (!strcmp(a, "--spread")) { o.spread = (float)atof(v); ++i; }
        else if (!strcmp(a, "--tone")) o.tone = 1;
        else if (!strcmp(a, "--srgb")) o.srgb = 1;
        else { fprintf(stderr, "unknown argument %s\n", a); return 2; }
    }

    mrt_scene* scene = NULL;
    size_t len = strlen(o.scene_name);
    if (len > 4 && !strcmp(o.scene_name + len - 4, ".obj")) CHECK(mrt_scene_load_obj(o.scene_name, &scene));
    else CHECK(mrt_scene_create_procedural(o.scene_name, o.seed, &scene));
    mrt_scene_view view;
    CHECK(mrt_scene_get(scene, &view));

    if (o.n_devices > 1) {
#ifdef MRT_WITH_NCCL
        if (o.resume || o.checkpoint || o.tone || o.srgb) { fprintf(stderr, "--resume/--checkpoint/--tone/--srgb need a single device\n"); return 2; }
        int rc = render_multi(&o, scene, &view);
        mrt_scene_destroy(scene);
        return rc;
#else
        fprintf(stderr, "--devices with several GPUs needs a build with NCCL (make: MRT_WITH_NCCL)\n");
        return 2;
#endif
    }

    unsigned w = o.w, h = o.h, frames = o.frames;
    mrt_config cfg;
    mrt_config_default(&cfg);
    cfg.device = o.devices[0];
    cfg.max_path_length = o.depth;
    cfg.pipeline = o.pipeline;
    mrt_ctx* ctx = NULL;
    CHECK(mrt_create(&cfg, &ctx));
    CHECK(mrt_set_scene_from(ctx, scene));
    if (o.optics) CHECK(mrt_set_extensions(ctx, o.optics, o.spread));
    CHECK(mrt_resize(ctx, w, h));

    if (o.resume) {                                              /* restores size, image, frame index and noise slots */
        CHECK(mrt_checkpoint_load(ctx, o.resume, &first));
        CHECK(mrt_get_size(ctx, &w, &h));
    }
    float* tiles = (float*)malloc((size_t)(frames ? frames : 1) * MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* init = (float*)malloc(MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* image = (float*)malloc((size_t)w * h * 4 * sizeof(float));
    if (!tiles || !init || !image) { fprintf(stderr, "out of memory\n"); return 1; }
    mrt_noise_fill(o.seed, 0xFFFFFFFFu, init);                   /* all three slots start identical, Renderer.mm:115-126 */
    for (unsigned s = 0; s < MRT_NOISE_SLOTS && !o.resume; ++s) CHECK(mrt_set_noise(ctx, s, init));
    for (unsigned f = 0; f < frames; ++f) mrt_noise_fill(o.seed, first + f, tiles + (size_t)f * MRT_NOISE_TILE_FLOATS);
    CHECK(mrt_sync(ctx));

    double t0 = now_s();
    CHECK(mrt_render_frames(ctx, first, frames, 1, tiles));
    CHECK(mrt_read_image(ctx, image));
    double t1 = now_s();

    mrt_stats st;
    CHECK(mrt_get_stats(ctx, &st));
    double samples = (double)w * h * frames;
    printf("{\"scene\": \"%s\", \"triangles\": %u, \"lights\": %u, \"width\": %u, \"height\": %u, \"frames\": %u, \"depth\": %u, "
           "\"device_ms\": %.3f, \"kernel_ms\": %.3f, \"wall_ms\": %.3f, \"msamples_per_s\": %.2f, \"mrays_per_s\": %.2f, \"primary_rays\": %llu, "
           "\"shadow_rays\": %llu, \"kernel_launches\": %llu, \"waves\": %u, \"bvh_nodes\": %u, \"bvh_depth\": %u, \"bvh_build_ms\": %.1f, "
           "\"bvh8_nodes\": %u, \"bvh8_depth\": %u, \"bvh8_build_ms\": %.1f, \"node_tests\": %llu, \"triangle_tests\": %llu, "
           "\"profile_trace_ms\": %.3f, \"profile_trace_launches\": %u}\n",
           o.scene_name, view.triangle_count, view.light_count, w, h, frames, o.depth, st.last_render_ms, st.last_kernel_ms, (t1 - t0) * 1e3,
           samples / (st.last_render_ms * 1e3), (double)(st.primary_rays + st.shadow_rays) / (st.last_render_ms * 1e3),
           (unsigned long long)st.primary_rays, (unsigned long long)st.shadow_rays, (unsigned long long)st.kernel_launches, st.waves_traced,
           st.bvh_nodes, st.bvh_depth, st.bvh_build_ms, st.bvh8_nodes, st.bvh8_depth, st.bvh8_build_ms,
           (unsigned long long)st.nodes_visited, (unsigned long long)st.triangles_tested,   /* MRT_TRACE_COUNT=1 fills these two */
           st.profile_trace_ms, st.profile_trace_launches);                                  /* MRT_WAVE_LOG=file fills these two */
    if (o.checkpoint) CHECK(mrt_checkpoint_save(ctx, o.checkpoint));
    if (o.out && (o.tone || o.srgb)) {
        mrt_display_params dp;
        mrt_display_params_default(&dp);
        dp.tone_mapping = o.tone; dp.manual_srgb = o.srgb;
        CHECK(mrt_display(ctx, &dp, NULL, 0, 0, image));
    }
    if (o.out && write_pfm(o.out, image, w, h)) { fprintf(stderr, "cannot write %s\n", o.out); return 1; }

    free(tiles); free(init); free(image);
    mrt_destroy(ctx);
    mrt_scene_destroy(scene);
    return 0;
}
