/*
 * mrt_render.c — headless plain-C driver over the C-ABI (include/mrt.h).
 *
 * Replaces the windowed frame loop of the reference (-drawInMTKView:, renderer/Renderer.mm:587-638, and the
 * NSApplication shells under macos/ and ios/): build or load a scene, render N frames, write the accumulated
 * image as a PFM file, print one JSON line with the reference's own statistic ("Mrays/s" = pixel-samples per
 * second, Renderer.mm:631-637) plus live ray segments per second.
 *
 * usage: mrt_render [--scene NAME|file.obj] [--size WxH] [--frames N] [--depth L] [--seed S] [--device D]
 *                   [--pipeline auto|wavefront|fused|staged] [--out image.pfm]
 *                   [--tone] [--srgb]            presentation pass on the written image (Shaders.metal:43-51)
 *                   [--resume file.ckpt]         continue an earlier render (frames then count from the stored index)
 *                   [--checkpoint file.ckpt]     save image + frame index + noise slots after rendering
 */
#define _POSIX_C_SOURCE 200809L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "mrt.h"

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

int main(int argc, char** argv) {
    const char* scene_name = "cornell";
    const char* out = NULL;
    const char* resume = NULL;
    const char* checkpoint = NULL;
    unsigned tone = 0, srgb = 0, first = 0;
    unsigned w = 256, h = 256, frames = 16, depth = MRT_MAX_PATH_LENGTH, device = 0, pipeline = MRT_PIPELINE_AUTO;
    unsigned long long seed = 1;
    for (int i = 1; i < argc; ++i) {
        const char* a = argv[i];
        const char* v = i + 1 < argc ? argv[i + 1] : "";
        if (!strcmp(a, "--scene")) { scene_name = v; ++i; }
        else if (!strcmp(a, "--size")) { if (sscanf(v, "%ux%u", &w, &h) != 2) { fprintf(stderr, "bad --size\n"); return 2; } ++i; }
        else if (!strcmp(a, "--frames")) { frames = (unsigned)atoi(v); ++i; }
        else if (!strcmp(a, "--depth")) { depth = (unsigned)atoi(v); ++i; }
        else if (!strcmp(a, "--seed")) { seed = strtoull(v, NULL, 10); ++i; }
        else if (!strcmp(a, "--device")) { device = (unsigned)atoi(v); ++i; }
        else if (!strcmp(a, "--pipeline")) {
            pipeline = !strcmp(v, "staged") ? MRT_PIPELINE_STAGED : !strcmp(v, "fused") ? MRT_PIPELINE_MEGAKERNEL
                     : !strcmp(v, "wavefront") ? MRT_PIPELINE_WAVEFRONT : MRT_PIPELINE_AUTO;
            ++i;
        }
        else if (!strcmp(a, "--out")) { out = v; ++i; }
        else if (!strcmp(a, "--resume")) { resume = v; ++i; }
        else if (!strcmp(a, "--checkpoint")) { checkpoint = v; ++i; }
        else if (!strcmp(a, "--tone")) tone = 1;
        else if (!strcmp(a, "--srgb")) srgb = 1;
        else { fprintf(stderr, "unknown argument %s\n", a); return 2; }
    }

    mrt_scene* scene = NULL;
    size_t len = strlen(scene_name);
    if (len > 4 && !strcmp(scene_name + len - 4, ".obj")) CHECK(mrt_scene_load_obj(scene_name, &scene));
    else CHECK(mrt_scene_create_procedural(scene_name, seed, &scene));
    mrt_scene_view view;
    CHECK(mrt_scene_get(scene, &view));

    mrt_config cfg;
    mrt_config_default(&cfg);
    cfg.device = (int)device;
    cfg.max_path_length = depth;
    cfg.pipeline = pipeline;
    mrt_ctx* ctx = NULL;
    CHECK(mrt_create(&cfg, &ctx));
    CHECK(mrt_set_scene_from(ctx, scene));
    CHECK(mrt_resize(ctx, w, h));

    if (resume) {                                                /* restores size, image, frame index and noise slots */
        CHECK(mrt_checkpoint_load(ctx, resume, &first));
        CHECK(mrt_get_size(ctx, &w, &h));
    }
    float* tiles = (float*)malloc((size_t)(frames ? frames : 1) * MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* init = (float*)malloc(MRT_NOISE_TILE_FLOATS * sizeof(float));
    float* image = (float*)malloc((size_t)w * h * 4 * sizeof(float));
    if (!tiles || !init || !image) { fprintf(stderr, "out of memory\n"); return 1; }
    mrt_noise_fill(seed, 0xFFFFFFFFu, init);                     /* all three slots start identical, Renderer.mm:115-126 */
    for (unsigned s = 0; s < MRT_NOISE_SLOTS && !resume; ++s) CHECK(mrt_set_noise(ctx, s, init));
    for (unsigned f = 0; f < frames; ++f) mrt_noise_fill(seed, first + f, tiles + (size_t)f * MRT_NOISE_TILE_FLOATS);
    CHECK(mrt_sync(ctx));

    double t0 = now_s();
    CHECK(mrt_render_frames(ctx, first, frames, 1, tiles));
    CHECK(mrt_read_image(ctx, image));
    double t1 = now_s();

    mrt_stats st;
    CHECK(mrt_get_stats(ctx, &st));
    double samples = (double)w * h * frames;
    printf("{\"scene\": \"%s\", \"triangles\": %u, \"lights\": %u, \"width\": %u, \"height\": %u, \"frames\": %u, \"depth\": %u, "
           "\"device_ms\": %.3f, \"wall_ms\": %.3f, \"msamples_per_s\": %.2f, \"mrays_per_s\": %.2f, \"primary_rays\": %llu, "
           "\"shadow_rays\": %llu, \"kernel_launches\": %llu, \"bvh_nodes\": %u, \"bvh_depth\": %u, \"bvh_build_ms\": %.1f, "
           "\"bvh8_nodes\": %u, \"bvh8_depth\": %u, \"node_tests\": %llu, \"triangle_tests\": %llu}\n",
           scene_name, view.triangle_count, view.light_count, w, h, frames, depth, st.last_render_ms, (t1 - t0) * 1e3,
           samples / (st.last_render_ms * 1e3), (double)(st.primary_rays + st.shadow_rays) / (st.last_render_ms * 1e3),
           (unsigned long long)st.primary_rays, (unsigned long long)st.shadow_rays, (unsigned long long)st.kernel_launches,
           st.bvh_nodes, st.bvh_depth, st.bvh_build_ms, st.bvh8_nodes, st.bvh8_depth,
           (unsigned long long)st.nodes_visited, (unsigned long long)st.triangles_tested);   /* MRT_TRACE_COUNT=1 fills the last two */
    if (checkpoint) CHECK(mrt_checkpoint_save(ctx, checkpoint));
    if (out && (tone || srgb)) {
        mrt_display_params dp;
        mrt_display_params_default(&dp);
        dp.tone_mapping = tone; dp.manual_srgb = srgb;
        CHECK(mrt_display(ctx, &dp, NULL, 0, 0, image));
    }
    if (out && write_pfm(out, image, w, h)) { fprintf(stderr, "cannot write %s\n", out); return 1; }

    free(tiles); free(init); free(image);
    mrt_destroy(ctx);
    mrt_scene_destroy(scene);
    return 0;
}
