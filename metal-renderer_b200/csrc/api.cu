// api.cu — the C-ABI of include/mrt.h: context, device memory, scene upload, stage entry points, fused render.
//
// Host-side counterpart of the reference's Renderer object (renderer/Renderer.mm): loadMetalWithView (:89-154),
// initRaytracing (:255-470), updateSharedData (:472-498), performRaytracing (:500-585),
// drawableSizeWillChange (:640-657).  No CPU fallback: every entry point that computes needs the CUDA device.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>
#include "bvh.h"
#include "bvh8.h"
#include <random>
#include "kernels.h"
#include "mrt.h"
#include "scene.h"
#include "ctx_view.h"
#include <nvtx3/nvToolsExt.h>       // header-only; ranges carry the reference's encoder labels (Renderer.mm:505, 527, 557, 574)

using namespace mrt;

struct NvtxRange { explicit NvtxRange(const char* name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } };

static thread_local std::string g_last_error;
static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_last_error = buf;
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(MRT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define NEED(cond, code, ...) do { if (!(cond)) return fail(code, __VA_ARGS__); } while (0)

#define MRT_STACK_DEPTH_HOST 62   /* traversal stack in device_trace.cuh holds 64 entries */
static const size_t kTileFloat4 = MRT_NOISE_DIMENSIONS * MRT_NOISE_DIMENSIONS;
static const size_t kTileBytes = kTileFloat4 * sizeof(float4);

struct mrt_ctx {
    mrt_config cfg{};
    cudaStream_t stream = nullptr;
    cudaStream_t groupStream[4] = {nullptr, nullptr, nullptr, nullptr};      // one per wavefront slot group
    cudaEvent_t evFork = nullptr, evJoin[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evK0 = nullptr, evK1 = nullptr, tableCopied = nullptr;
    cudaEvent_t user[8] = {};
    bool kernelTimed = false;
    bool timed = false;
    // scene
    DeviceScene S{};
    void* dNodes = nullptr; void* dTris = nullptr; void* dShade = nullptr; void* dMats = nullptr; void* dLights = nullptr;
    void* dCdf = nullptr;
    bool haveScene = false;
    // per-size buffers (Renderer.mm:645-653)
    uint32_t W = 0, H = 0, rowFirst = 0, rowStride = 1;
    mrt_ray* dRays = nullptr; mrt_light_sampling_ray* dLightRays = nullptr; mrt_intersection* dIsects = nullptr;
    float4* dImage = nullptr;
    // noise: pool tiles 0..2 are the three slots; tiles 3.. hold a fused batch
    float4* dNoisePool = nullptr; size_t poolTiles = 0;
    uint4* dFrameTable = nullptr; FrameEntry* hFrameTable = nullptr; size_t tableCap = 0;
    unsigned long long* dCounters = nullptr;     // primary, shadow, nodes, tris
    unsigned int* dHead = nullptr;               // work counter of the persistent intersector (staged entry points)
    void* dTriSlot = nullptr;                    // original triangle index -> leaf-order record
    void* dNodes8 = nullptr; void* dTris8 = nullptr;   // 8-wide tree (bvh8.h)
    WfBuffers wf{};                              // wavefront path state + queues (sized by mrt_resize)
    uint32_t wfSlotMult = 1;                     // path slots per pixel (MRT_WF_SLOTS)
    uint32_t nextFrame = 0;                      // frame index after the last one accumulated (checkpoint)
    bool profile = getenv("MRT_WAVE_LOG") != nullptr;                        // event pair around every trace launch
    bool countVisits = getenv("MRT_TRACE_COUNT") && atoi(getenv("MRT_TRACE_COUNT")) != 0;   // node / triangle tests into stats
    std::vector<cudaEvent_t> profEvents; size_t profUsed = 0;
    unsigned int* dWaveCounts = nullptr; size_t waveCountCap = 0;   // per-wave queue sizes (profile mode)
    float4* dSamples = nullptr; size_t sampleCap = 0;            // per-sample radiance of a wavefront batch
    void* wfArena = nullptr;
    float sceneLo[3] = {0, 0, 0}, sceneHi[3] = {1, 1, 1};        // bounding box of the scene's vertices (ray-ordering grid)
    // wave loop as a CUDA graph: first shade -> WHILE { trace, shade, trace, shade } -> resolve; rebuilt when its kernel
    // arguments change (another size, frame count, shard, accumulate mode ...), replayed otherwise
    cudaGraphExec_t wfGraph = nullptr; std::string wfGraphKey;
    cudaStream_t captureStream = nullptr;                        // body of the WHILE node is captured on this one
    uint32_t waveSchedule = MRT_SCHEDULE_GRAPH;                  // mrt_set_wave_schedule
    volatile unsigned int* hStop = nullptr; uint32_t batchSerial = 0;   // polling host loop: mapped word the trace kernel writes
    cudaEvent_t pollEvents[4] = {};                              // polling host loop: one event every 8 waves
    uint64_t graphKernels = 0; uint32_t kernelsPerWave = 2;      // kernels launched by graph replays (read back in mrt_get_stats)
    mrt_stats stats{};
    void* peer = nullptr;                                        // csrc/peer.cu (mrt_peer_*): exchange block + peer mappings

    Grid2D grid() const { Grid2D g; g.W = W; g.H = H; g.rowFirst = rowFirst; g.rowStride = rowStride; return g; }
    uint32_t ownedRows() const { return rowFirst < H ? (H - rowFirst + rowStride - 1) / rowStride : 0; }
};

static void free_scene(mrt_ctx* c) {
    cudaFree(c->dNodes); cudaFree(c->dTris); cudaFree(c->dShade); cudaFree(c->dMats); cudaFree(c->dLights); cudaFree(c->dCdf);
    cudaFree(c->dTriSlot); cudaFree(c->dNodes8);          // dTris8 lives in the same allocation
    c->dNodes8 = c->dTris8 = nullptr;
    c->dNodes = c->dTris = c->dShade = c->dMats = c->dLights = c->dCdf = c->dTriSlot = nullptr;
    c->wf.triSlot = nullptr; c->wf.triRecords = nullptr;
    c->haveScene = false;
}
static void free_sized(mrt_ctx* c) {
    peer_release(&c->peer);                                      // peers map dImage: the group dies with it
    cudaFree(c->dRays); cudaFree(c->dLightRays); cudaFree(c->dIsects); cudaFree(c->dImage); cudaFree(c->wfArena);
    c->dRays = nullptr; c->dLightRays = nullptr; c->dIsects = nullptr; c->dImage = nullptr; c->wfArena = nullptr;
    const uint32_t* triSlot = c->wf.triSlot; const float4* triRecords = c->wf.triRecords;
    c->wf = WfBuffers{};
    c->wf.triSlot = triSlot; c->wf.triRecords = triRecords;
}
static int ensure_pool(mrt_ctx* c, size_t tiles) {
    if (tiles <= c->poolTiles) return MRT_OK;
    float4* p = nullptr;
    CK(cudaMalloc(&p, tiles * kTileBytes));
    if (c->dNoisePool) {
        CK(cudaMemcpyAsync(p, c->dNoisePool, MRT_NOISE_SLOTS * kTileBytes, cudaMemcpyDeviceToDevice, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        cudaFree(c->dNoisePool);
    } else {
        CK(cudaMemsetAsync(p, 0, MRT_NOISE_SLOTS * kTileBytes, c->stream));
    }
    c->dNoisePool = p; c->poolTiles = tiles;
    return MRT_OK;
}
static int ensure_table(mrt_ctx* c, size_t entries) {
    if (entries <= c->tableCap) return MRT_OK;
    CK(cudaStreamSynchronize(c->stream));
    if (c->dFrameTable) cudaFree(c->dFrameTable);
    if (c->hFrameTable) cudaFreeHost(c->hFrameTable);
    c->dFrameTable = nullptr; c->hFrameTable = nullptr; c->tableCap = 0;
    CK(cudaMalloc(&c->dFrameTable, entries * sizeof(uint4)));
    CK(cudaMallocHost(&c->hFrameTable, entries * sizeof(FrameEntry)));
    c->tableCap = entries;
    return MRT_OK;
}

static int create_resources(mrt_ctx* c) {
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->captureStream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&c->ev0)); CK(cudaEventCreate(&c->ev1)); CK(cudaEventCreate(&c->evK0)); CK(cudaEventCreate(&c->evK1));
    CK(cudaEventCreateWithFlags(&c->evFork, cudaEventDisableTiming));
    for (int g = 0; g < 4; ++g) { CK(cudaStreamCreateWithFlags(&c->groupStream[g], cudaStreamNonBlocking)); CK(cudaEventCreateWithFlags(&c->evJoin[g], cudaEventDisableTiming)); }
    for (auto& e : c->user) CK(cudaEventCreate(&e));
    for (auto& e : c->pollEvents) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->tableCopied, cudaEventDisableTiming));
    CK(cudaMalloc(&c->dCounters, 4 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(c->dCounters, 0, 4 * sizeof(unsigned long long), c->stream));
    CK(cudaMalloc(&c->dHead, sizeof(unsigned int)));
    {   // mapped host word the trace kernel writes when a batch is over (polling host loop)
        void* h = nullptr;
        CK(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
        std::memset(h, 0, 64);
        c->hStop = static_cast<volatile unsigned int*>(h);
    }
    return ensure_pool(c, MRT_NOISE_SLOTS + 16);
}

extern "C" {

const char* mrt_last_error(void) { return g_last_error.c_str(); }
int mrt_version(void) { return MRT_VERSION; }

void mrt_config_default(mrt_config* cfg) {
    if (!cfg) return;
    std::memset(cfg, 0, sizeof *cfg);
    cfg->struct_size = sizeof(mrt_config);
    cfg->device = 0;
    cfg->max_path_length = MRT_MAX_PATH_LENGTH;      // Raytracing.h:23
    cfg->accumulate_mode = MRT_ACCUM_MEAN;           // ACCUMULATE_IMAGE true, Raytracing.h:14
    cfg->animate_noise = 1;                          // Raytracing.h:20
    cfg->pipeline = MRT_PIPELINE_AUTO;
    cfg->bvh_max_leaf = 4;
    cfg->bvh_threads = 0;
    cfg->count_rays = 1;
}

int mrt_create(const mrt_config* cfg, mrt_ctx** out) {
    NEED(out, MRT_E_INVALID, "mrt_create: out is NULL");
    *out = nullptr;
    mrt_config def; mrt_config_default(&def);
    if (cfg) { NEED(cfg->struct_size == sizeof(mrt_config), MRT_E_INVALID, "mrt_create: struct_size %u != %zu", cfg->struct_size, sizeof(mrt_config)); def = *cfg; }
    NEED(def.max_path_length >= 1 && def.max_path_length <= 64, MRT_E_INVALID, "max_path_length must be in 1..64");
    NEED(def.accumulate_mode <= MRT_ACCUM_REPLACE, MRT_E_INVALID, "bad accumulate_mode");
    NEED(def.pipeline <= MRT_PIPELINE_STAGED, MRT_E_INVALID, "bad pipeline");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(MRT_E_CUDA, "no CUDA device (%s); this library has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "count 0");
    NEED(def.device >= 0 && def.device < n, MRT_E_INVALID, "device %d out of range (0..%d)", def.device, n - 1);
    CK(cudaSetDevice(def.device));
    mrt_ctx* c = new (std::nothrow) mrt_ctx();
    NEED(c, MRT_E_NOMEM, "out of host memory");
    c->cfg = def;
    if (c->cfg.bvh_max_leaf == 0) c->cfg.bvh_max_leaf = 4;
    int rc = create_resources(c);
    if (rc != MRT_OK) { mrt_destroy(c); return rc; }       // a half-built context releases whatever it already holds
    *out = c;
    return MRT_OK;
}

int mrt_destroy(mrt_ctx* c) {
    if (!c) return MRT_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->wfGraph) cudaGraphExecDestroy(c->wfGraph);
    free_scene(c); free_sized(c);
    cudaFree(c->dNoisePool); cudaFree(c->dFrameTable); cudaFree(c->dCounters); cudaFree(c->dHead); cudaFree(c->dWaveCounts); cudaFree(c->dSamples);
    for (auto& e : c->profEvents) cudaEventDestroy(e);
    if (c->hFrameTable) cudaFreeHost(c->hFrameTable);
    for (cudaEvent_t e : {c->ev0, c->ev1, c->evK0, c->evK1, c->tableCopied}) if (e) cudaEventDestroy(e);
    for (auto& e : c->user) if (e) cudaEventDestroy(e);
    for (auto& e : c->pollEvents) if (e) cudaEventDestroy(e);
    if (c->hStop) cudaFreeHost(const_cast<unsigned int*>(c->hStop));
    if (c->captureStream) cudaStreamDestroy(c->captureStream);
    for (int g = 0; g < 4; ++g) { if (c->groupStream[g]) { cudaStreamSynchronize(c->groupStream[g]); cudaStreamDestroy(c->groupStream[g]); } if (c->evJoin[g]) cudaEventDestroy(c->evJoin[g]); }
    if (c->evFork) cudaEventDestroy(c->evFork);
    if (c->stream) cudaStreamDestroy(c->stream);
    cudaGetLastError();
    delete c;
    return MRT_OK;
}

// MRT_L2_PERSIST=1: keep the wide tree (and as much of its triangle records as fits) in the persisting part of L2, so
// that the shade kernel's streaming between two trace launches does not evict it.  Window = [base, base + bytes)
// clipped to the device's maximum; hit ratio = set-aside / window when the window is larger than the set-aside.
static cudaError_t set_l2_window(mrt_ctx* c, void* base, size_t bytes) {
    static const int on = [] { const char* e = getenv("MRT_L2_PERSIST"); return e ? atoi(e) : 0; }();
    if (!on || bytes == 0) return cudaSuccess;
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, c->cfg.device);
    if (e != cudaSuccess) return e;
    if (prop.persistingL2CacheMaxSize <= 0 || prop.accessPolicyMaxWindowSize <= 0) return cudaSuccess;
    size_t setAside = std::min<size_t>((size_t)prop.persistingL2CacheMaxSize, bytes);
    if (on > 1) setAside = std::min<size_t>(setAside, (size_t)on << 20);          // MRT_L2_PERSIST=<MiB> caps the set-aside
    e = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setAside);
    if (e != cudaSuccess) return e;
    cudaStreamAttrValue v;
    std::memset(&v, 0, sizeof v);
    v.accessPolicyWindow.base_ptr = base;
    v.accessPolicyWindow.num_bytes = std::min<size_t>(bytes, (size_t)prop.accessPolicyMaxWindowSize);
    v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)setAside / (double)v.accessPolicyWindow.num_bytes);
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    e = cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &v);
    for (int g = 0; g < 4 && e == cudaSuccess; ++g) e = cudaStreamSetAttribute(c->groupStream[g], cudaStreamAttributeAccessPolicyWindow, &v);
    if (getenv("MRT_L2_VERBOSE"))
        fprintf(stderr, "mrt: L2 window %zu B, set-aside %zu B (device max %d B, max window %d B), hit ratio %.3f\n",
                (size_t)v.accessPolicyWindow.num_bytes, setAside, prop.persistingL2CacheMaxSize, prop.accessPolicyMaxWindowSize,
                v.accessPolicyWindow.hitRatio);
    return e;
}

// Host half of -initRaytracing (Renderer.mm:450-469): validation, both trees, records repacked for 128-bit loads.
// Built once, uploaded to any number of contexts (one per GPU) by mrt_set_prepared_scene.
// Walker/Vose alias table of n probabilities (optics extension, MRT_EXT_ALIAS_LIGHTS): slot i keeps itself with
// probability prob[i] and hands the rest to alias[i], so that P(i) = (prob[i] + sum over slots j with alias[j] == i of
// (1 - prob[j])) / n = pdf[i] / sum(pdf).  Worked in double, small / large work lists in index order: deterministic.
static void build_alias_table(const float* pdf, uint32_t n, float* prob, uint32_t* alias) {
    if (n == 0) return;
    double total = 0.0;
    for (uint32_t i = 0; i < n; ++i) total += pdf[i] > 0.0f ? (double)pdf[i] : 0.0;
    std::vector<double> q(n);
    std::vector<uint32_t> small, large;
    for (uint32_t i = 0; i < n; ++i) {
        q[i] = total > 0.0 ? (pdf[i] > 0.0f ? (double)pdf[i] : 0.0) * n / total : 1.0;
        alias[i] = i;
        (q[i] < 1.0 ? small : large).push_back(i);
    }
    size_t si = 0, li = 0;
    while (si < small.size() && li < large.size()) {
        const uint32_t sm = small[si++], lg = large[li];
        prob[sm] = (float)q[sm]; alias[sm] = lg;
        q[lg] -= 1.0 - q[sm];
        if (q[lg] < 1.0) { small.push_back(lg); ++li; }
    }
    for (; li < large.size(); ++li) prob[large[li]] = 1.0f;
    for (; si < small.size(); ++si) prob[small[si]] = 1.0f;        // rounding leftovers
}

struct mrt_prepared_scene {
    BvhResult bvh; Bvh8Result bvh8;
    std::vector<float> shade, mats, lts;
    std::vector<uint32_t> cdfAndGuide, triSlotWide, triSlotBinary;
    uint32_t cdfCount = 0, guideBuckets = 0, aliasOffset = 0, lightCount = 0, triangleCount = 0;
    float lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
};

int mrt_scene_prepare(const mrt_vertex* vertices, uint32_t vertex_count, const uint32_t* indices, uint32_t triangle_count,
                      const mrt_triangle_reference* references, const mrt_material* materials, uint32_t material_count,
                      const mrt_light_triangle* lights, uint32_t light_count, uint32_t bvh_max_leaf, uint32_t bvh_threads,
                      mrt_prepared_scene** out) {
    NEED(out, MRT_E_INVALID, "out is NULL");
    *out = nullptr;
    NEED(vertices && indices && references && materials && lights, MRT_E_INVALID, "mrt_set_scene: NULL array");
    NEED(triangle_count < (1u << 28), MRT_E_INVALID, "triangle_count must be < 2^28");
    for (size_t i = 0; i < 3 * (size_t)triangle_count; ++i)
        NEED(indices[i] < vertex_count, MRT_E_INVALID, "index %zu = %u out of range (vertex_count %u)", i, indices[i], vertex_count);
    for (uint32_t t = 0; t < triangle_count; ++t) {
        const mrt_triangle_reference& r = references[t];
        NEED(r.tri[0] < vertex_count && r.tri[1] < vertex_count && r.tri[2] < vertex_count, MRT_E_INVALID, "reference %u: vertex index out of range", t);
        NEED(r.materialIndex < material_count, MRT_E_INVALID, "reference %u: materialIndex %u out of range", t, r.materialIndex);
        NEED(r.lightTriangleIndex == MRT_NOT_A_LIGHT || r.lightTriangleIndex < light_count, MRT_E_INVALID, "reference %u: lightTriangleIndex out of range", t);
    }
    for (uint32_t l = 0; l < light_count; ++l) NEED(lights[l].index < triangle_count, MRT_E_INVALID, "light %u: index out of range", l);
    // selectLightTriangle (KernelHelpers.h:49-54) scans the cdf linearly; the kernels search it by bisection from a guide
    // table, which selects the same light only if the cdf does not decrease (Renderer.mm:435-448 builds it that way)
    for (uint32_t l = 0; l < light_count; ++l)
        NEED(lights[l].cdf <= lights[l + 1].cdf, MRT_E_INVALID, "light cdf decreases at %u (%g > %g): build it as Renderer.mm:435-448 does", l,
             (double)lights[l].cdf, (double)lights[l + 1].cdf);
    mrt_prepared_scene* P = new (std::nothrow) mrt_prepared_scene();
    NEED(P, MRT_E_NOMEM, "out of host memory");
    if (bvh_max_leaf == 0) bvh_max_leaf = 4;
    {   // the binary tree (megakernel, mrt_intersect_host) is built beside the wide one: their serial parts overlap
        std::thread binary([&] { build_bvh(vertices, indices, triangle_count, bvh_max_leaf, bvh_threads, P->bvh); });
        build_bvh8(vertices, indices, triangle_count, bvh_threads, P->bvh8);
        binary.join();
    }
    if (!(P->bvh.depth < MRT_STACK_DEPTH_HOST)) { const uint32_t d = P->bvh.depth; delete P; return fail(MRT_E_INVALID, "BVH depth %u exceeds traversal stack", d); }
    if (!(P->bvh8.depth + 2 < 32)) { const uint32_t d = P->bvh8.depth; delete P; return fail(MRT_E_INVALID, "wide BVH depth %u exceeds traversal stack", d); }

    // shading records: vertices gathered per triangle through TriangleReference.tri (Shaders.metal:129,140)
    P->shade.assign(20 * (size_t)std::max(1u, triangle_count), 0.0f);
    for (uint32_t t = 0; t < triangle_count; ++t) {
        float* s = &P->shade[20 * (size_t)t];
        const mrt_triangle_reference& r = references[t];
        for (int k = 0; k < 3; ++k) { std::memcpy(s + 3 * k, vertices[r.tri[k]].v, 12); std::memcpy(s + 9 + 3 * k, vertices[r.tri[k]].n, 12); }
        std::memcpy(s + 18, &r.materialIndex, 4); std::memcpy(s + 19, &r.lightTriangleIndex, 4);
    }
    P->mats.assign(8 * (size_t)std::max(1u, material_count), 0.0f);
    for (uint32_t m = 0; m < material_count; ++m) std::memcpy(&P->mats[8 * (size_t)m], &materials[m], 32);
    P->lts.assign(24 * (size_t)(light_count + 1), 0.0f);
    std::vector<float> cdf(light_count + 1);
    for (uint32_t l = 0; l <= light_count; ++l) {
        const mrt_light_triangle& L = lights[l];
        float* p = &P->lts[24 * (size_t)l];
        std::memcpy(p, L.emissive.values, 12); p[3] = L.area;
        std::memcpy(p + 4, L.v1.v, 12); p[7] = L.pdf;
        std::memcpy(p + 8, L.v1.n, 12); std::memcpy(p + 11, &L.index, 4);
        std::memcpy(p + 12, L.v2.v, 12); std::memcpy(p + 15, L.v2.n, 12);
        std::memcpy(p + 18, L.v3.v, 12); std::memcpy(p + 21, L.v3.n, 12);
        cdf[l] = L.cdf;
    }
    // guide table for selectLightTriangle (device_shading.cuh): the exact host search at the start of every slice
    uint32_t buckets = 0;
    std::vector<uint32_t> guide;
    if (light_count >= 32) {
        buckets = 64; while (buckets < 4u * light_count && buckets < 16384u) buckets <<= 1;
        guide.resize(buckets + 1);
        auto select = [&](float xi) { uint32_t i = 0; for (; i < light_count && cdf[i + 1] <= xi; ++i) {} return i; };   // KernelHelpers.h:49-54
        for (uint32_t b = 0; b < buckets; ++b) guide[b] = select(float(b) / float(buckets));
        guide[buckets] = light_count;
    }
    P->cdfAndGuide.resize(cdf.size() + guide.size());
    std::memcpy(P->cdfAndGuide.data(), cdf.data(), cdf.size() * 4);
    if (!guide.empty()) std::memcpy(P->cdfAndGuide.data() + cdf.size(), guide.data(), guide.size() * 4);
    P->cdfCount = (uint32_t)cdf.size(); P->guideBuckets = buckets;
    {   // alias table of the same pdfs, 8-byte aligned behind cdf and guide: (prob, alias) per light
        P->aliasOffset = (uint32_t)((P->cdfAndGuide.size() + 1u) & ~size_t(1));
        P->cdfAndGuide.resize(P->aliasOffset + 2 * (size_t)light_count, 0u);
        std::vector<float> pdf(light_count), prob(light_count);
        std::vector<uint32_t> alias(light_count);
        for (uint32_t l = 0; l < light_count; ++l) pdf[l] = lights[l].pdf;
        build_alias_table(pdf.data(), light_count, prob.data(), alias.data());
        for (uint32_t l = 0; l < light_count; ++l) {
            std::memcpy(&P->cdfAndGuide[P->aliasOffset + 2 * (size_t)l], &prob[l], 4);
            P->cdfAndGuide[P->aliasOffset + 2 * (size_t)l + 1] = alias[l];
        }
    }
    P->triSlotWide.assign(std::max(1u, triangle_count), 0u);
    for (size_t k = 0; k < P->bvh8.tris.size(); ++k) P->triSlotWide[P->bvh8.tris[k].index] = (uint32_t)k;
    P->triSlotBinary.assign(std::max(1u, triangle_count), 0u);
    for (size_t k = 0; k < P->bvh.tris.size(); ++k) P->triSlotBinary[P->bvh.tris[k].index] = (uint32_t)k;
    P->lightCount = light_count; P->triangleCount = triangle_count;
    if (vertex_count) {
        for (int a = 0; a < 3; ++a) { P->lo[a] = INFINITY; P->hi[a] = -INFINITY; }
        for (uint32_t v = 0; v < vertex_count; ++v)
            for (int a = 0; a < 3; ++a) { const float x = vertices[v].v[a]; P->lo[a] = std::min(P->lo[a], x); P->hi[a] = std::max(P->hi[a], x); }
    }
    *out = P;
    return MRT_OK;
}
void mrt_prepared_scene_destroy(mrt_prepared_scene* p) { delete p; }

// Device half: upload to THIS context.  The prepared scene is only read, so several contexts (threads) may share it.
int mrt_set_prepared_scene(mrt_ctx* c, const mrt_prepared_scene* P) {
    NEED(c && P, MRT_E_INVALID, "NULL argument");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    free_scene(c);
    if (c->wfGraph) { cudaGraphExecDestroy(c->wfGraph); c->wfGraph = nullptr; c->wfGraphKey.clear(); }
    const BvhResult& bvh = P->bvh; const Bvh8Result& bvh8 = P->bvh8;
    auto upload = [&](void** dst, const void* src, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, std::max<size_t>(bytes, 16));
        if (e != cudaSuccess) return e;
        return bytes ? cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
    };
    CK(upload(&c->dNodes, bvh.nodes.data(), bvh.nodes.size() * sizeof(BvhNode)));
    CK(upload(&c->dTris, bvh.tris.data(), bvh.tris.size() * sizeof(BvhTri)));
    CK(upload(&c->dShade, P->shade.data(), P->shade.size() * 4));
    CK(upload(&c->dMats, P->mats.data(), P->mats.size() * 4));
    CK(upload(&c->dLights, P->lts.data(), P->lts.size() * 4));
    CK(upload(&c->dCdf, P->cdfAndGuide.data(), P->cdfAndGuide.size() * 4));
    c->S.lightGuide = P->guideBuckets ? reinterpret_cast<const uint32_t*>(c->dCdf) + P->cdfCount : nullptr;
    c->S.lightGuideBuckets = P->guideBuckets;
    c->S.lightAlias = reinterpret_cast<const float2*>(reinterpret_cast<const uint32_t*>(c->dCdf) + P->aliasOffset);
    {   // wide nodes and their triangle records share ONE allocation (nodes first), so that a single L2 access-policy
        // window can cover what the trace kernel gathers from
        const size_t nodeBytes = (bvh8.nodes.size() * sizeof(Bvh8Node) + 255) & ~size_t(255);
        const size_t triBytes = bvh8.tris.size() * sizeof(BvhTri);
        CK(cudaMalloc(&c->dNodes8, std::max<size_t>(nodeBytes + triBytes, 16)));
        c->dTris8 = static_cast<char*>(c->dNodes8) + nodeBytes;
        if (!bvh8.nodes.empty()) CK(cudaMemcpy(c->dNodes8, bvh8.nodes.data(), bvh8.nodes.size() * sizeof(Bvh8Node), cudaMemcpyHostToDevice));
        if (triBytes) CK(cudaMemcpy(c->dTris8, bvh8.tris.data(), triBytes, cudaMemcpyHostToDevice));
        CK(set_l2_window(c, c->dNodes8, nodeBytes + triBytes));
    }
    c->S.nodes8 = (const float4*)c->dNodes8; c->S.tris8 = (const float4*)c->dTris8;
    const bool wide = trace_uses_wide_bvh();
    const std::vector<uint32_t>& triSlot = wide ? P->triSlotWide : P->triSlotBinary;
    CK(upload(&c->dTriSlot, triSlot.data(), triSlot.size() * 4));
    c->wf.triSlot = (const uint32_t*)c->dTriSlot;
    c->wf.triRecords = (const float4*)(wide ? c->dTris8 : c->dTris);
    c->stats.bvh8_nodes = (uint32_t)bvh8.nodes.size(); c->stats.bvh8_depth = bvh8.depth;
    c->S.nodes = (const float4*)c->dNodes; c->S.tris = (const float4*)c->dTris; c->S.shadeTris = (const float4*)c->dShade;
    c->S.materials = (const float4*)c->dMats; c->S.lights = (const float4*)c->dLights; c->S.lightCdf = (const float*)c->dCdf;
    c->S.lightCount = P->lightCount; c->S.triangleCount = P->triangleCount;
    c->stats.bvh_nodes = (uint32_t)bvh.nodes.size(); c->stats.bvh_depth = bvh.depth; c->stats.bvh_sah_cost = bvh.sah_cost;
    c->stats.bvh_build_ms = (float)bvh.build_ms;
    c->stats.bvh8_build_ms = (float)bvh8.build_ms;
    for (int a = 0; a < 3; ++a) { c->sceneLo[a] = P->lo[a]; c->sceneHi[a] = P->hi[a]; }
    c->haveScene = true;
    return MRT_OK;
}

int mrt_set_scene(mrt_ctx* c, const mrt_vertex* vertices, uint32_t vertex_count, const uint32_t* indices, uint32_t triangle_count,
                  const mrt_triangle_reference* references, const mrt_material* materials, uint32_t material_count,
                  const mrt_light_triangle* lights, uint32_t light_count) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    mrt_prepared_scene* P = nullptr;
    int rc = mrt_scene_prepare(vertices, vertex_count, indices, triangle_count, references, materials, material_count, lights, light_count,
                               c->cfg.bvh_max_leaf, c->cfg.bvh_threads, &P);
    if (rc != MRT_OK) return rc;
    rc = mrt_set_prepared_scene(c, P);
    delete P;
    return rc;
}

int mrt_scene_prepare_from(const mrt_scene* s, uint32_t bvh_threads, mrt_prepared_scene** out) {
    NEED(s && out, MRT_E_INVALID, "NULL argument");
    return mrt_scene_prepare(s->vertices.data(), (uint32_t)s->vertices.size(), s->indices.data(), (uint32_t)s->references.size(),
                             s->references.data(), s->materials.data(), (uint32_t)s->materials.size(), s->lights.data(), s->light_count,
                             4, bvh_threads, out);
}

int mrt_set_scene_from(mrt_ctx* c, const mrt_scene* s) {
    NEED(c && s, MRT_E_INVALID, "NULL argument");
    return mrt_set_scene(c, s->vertices.data(), (uint32_t)s->vertices.size(), s->indices.data(), (uint32_t)s->references.size(),
                         s->references.data(), s->materials.data(), (uint32_t)s->materials.size(), s->lights.data(), s->light_count);
}

int mrt_resize(mrt_ctx* c, uint32_t width, uint32_t height) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    NEED(width >= 2 && height >= 2, MRT_E_INVALID, "width and height must be >= 2 (raygen divides by size-1, Shaders.metal:92-94)");
    NEED((uint64_t)width * height < (1ull << 31), MRT_E_INVALID, "too many rays");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    free_sized(c);
    size_t n = (size_t)width * height;
    CK(cudaMalloc(&c->dRays, n * sizeof(mrt_ray)));
    CK(cudaMalloc(&c->dLightRays, n * sizeof(mrt_light_sampling_ray)));
    CK(cudaMalloc(&c->dIsects, n * sizeof(mrt_intersection)));
    CK(cudaMalloc(&c->dImage, n * sizeof(float4)));
    CK(cudaMemsetAsync(c->dRays, 0, n * sizeof(mrt_ray), c->stream));                    // new MTLBuffers are zero-filled
    CK(cudaMemsetAsync(c->dLightRays, 0, n * sizeof(mrt_light_sampling_ray), c->stream));
    CK(cudaMemsetAsync(c->dIsects, 0, n * sizeof(mrt_intersection), c->stream));
    CK(cudaMemsetAsync(c->dImage, 0, n * sizeof(float4), c->stream));
    c->W = width; c->H = height;
    if (c->wfGraph) { cudaGraphExecDestroy(c->wfGraph); c->wfGraph = nullptr; c->wfGraphKey.clear(); }
    {   // wavefront path state: one slot per pixel of an 8x4-tiled grid, SoA float4 arrays + two queue pairs
        // MRT_WF_SLOTS: path slots per pixel (default 1).  More slots than pixels = fewer, larger waves.
#ifdef MRT_EXPERIMENTS
        static const int slotMult = std::min(4, std::max(1, getenv("MRT_WF_SLOTS") ? atoi(getenv("MRT_WF_SLOTS")) : 1));
#else
        const int slotMult = 1;
#endif
        c->wfSlotMult = (uint32_t)slotMult;
        const size_t slots = (size_t)((width + 7u) / 8u) * ((height + 3u) / 4u) * 32u * slotMult;
        const size_t f4 = slots * sizeof(float4), u4 = slots * sizeof(uint32_t);
        const size_t regions = slots / 32;                                  // smallest region = one 8x4 tile
        const size_t histBytes = 4 * 4 * (size_t)wf_sort_buckets(3) * sizeof(uint32_t);   // up to 4 groups x (2 counts + 2 cursors) x 4096 buckets
        const size_t ctlBytes = 128 + MRT_CTL_WORDS * sizeof(uint32_t) + 96;             // 128 = 4 groups x (counts[4], heads[2], 2 spare)
        const size_t total = 7 * f4 + 10 * u4 + ctlBytes + histBytes + 2 * regions * sizeof(uint32_t);
        CK(cudaMalloc(&c->wfArena, total));
        CK(cudaMemsetAsync(c->wfArena, 0, total, c->stream));
        char* p = static_cast<char*>(c->wfArena);
        auto take4 = [&]() { float4* r = reinterpret_cast<float4*>(p); p += f4; return r; };
        auto takeU = [&]() { uint32_t* r = reinterpret_cast<uint32_t*>(p); p += u4; return r; };
        c->wf.P0 = take4(); c->wf.P1 = take4(); c->wf.P2 = take4(); c->wf.P3 = take4(); c->wf.H = take4();
        c->wf.S0 = take4(); c->wf.S1 = take4();
        c->wf.visible = takeU(); c->wf.item = takeU();
        c->wf.queueExt[0] = takeU(); c->wf.queueExt[1] = takeU(); c->wf.queueShadow[0] = takeU(); c->wf.queueShadow[1] = takeU();
        for (int k = 0; k < 2; ++k) { c->wf.sortSlot[k] = takeU(); c->wf.sortKey[k] = takeU(); }
        c->wf.counts = reinterpret_cast<unsigned int*>(p); c->wf.heads = c->wf.counts + 4; p += 128;
        c->wf.ctl = reinterpret_cast<unsigned int*>(p); p += MRT_CTL_WORDS * sizeof(uint32_t) + 96;
        c->wf.sortHist = reinterpret_cast<unsigned int*>(p); p += histBytes;
        c->wf.regionNext = reinterpret_cast<unsigned int*>(p); c->wf.blockRegion = c->wf.regionNext + regions;
        c->wf.slots = (uint32_t)slots; c->wf.pixels = (uint32_t)slots;
    }
    return MRT_OK;
}

int mrt_get_size(mrt_ctx* c, uint32_t* width, uint32_t* height) {
    NEED(c && width && height, MRT_E_INVALID, "NULL argument");
    *width = c->dImage ? c->W : 0; *height = c->dImage ? c->H : 0;
    return MRT_OK;
}

int mrt_set_noise(mrt_ctx* c, uint32_t slot, const float* tile) {
    NEED(c && tile, MRT_E_INVALID, "NULL argument");
    NEED(slot < MRT_NOISE_SLOTS, MRT_E_INVALID, "slot %u out of range", slot);
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(c->dNoisePool + slot * kTileFloat4, tile, kTileBytes, cudaMemcpyDefault, c->stream));   // host or device tile
    return MRT_OK;
}

int mrt_set_row_shard(mrt_ctx* c, uint32_t first_row, uint32_t row_stride) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    NEED(row_stride >= 1 && first_row < row_stride, MRT_E_INVALID, "need row_stride >= 1 and first_row < row_stride");
    c->rowFirst = first_row; c->rowStride = row_stride;
    return MRT_OK;
}
int mrt_set_accumulate_mode(mrt_ctx* c, uint32_t mode) {
    NEED(c && mode <= MRT_ACCUM_REPLACE, MRT_E_INVALID, "bad accumulate mode");
    c->cfg.accumulate_mode = mode;
    return MRT_OK;
}
int mrt_set_profile(mrt_ctx* c, uint32_t on) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    c->profile = (on & MRT_PROFILE_TRACE_TIME) != 0;
    c->countVisits = (on & MRT_PROFILE_VISITS) != 0;
    return MRT_OK;
}
int mrt_set_wave_schedule(mrt_ctx* c, uint32_t schedule) {
    NEED(c && schedule <= MRT_SCHEDULE_HOST_LOOP, MRT_E_INVALID, "bad wave schedule");
    c->waveSchedule = schedule;
    return MRT_OK;
}
int mrt_set_extensions(mrt_ctx* c, uint32_t flags, float dispersion) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    NEED((flags & ~(uint32_t)(MRT_EXT_REFRACTION | MRT_EXT_DISPERSION | MRT_EXT_ALIAS_LIGHTS | MRT_EXT_CONTINUOUS_SPECTRUM)) == 0, MRT_E_INVALID,
         "unknown extension bit in 0x%x", flags);
    NEED(!(flags & MRT_EXT_DISPERSION) || (flags & MRT_EXT_REFRACTION), MRT_E_INVALID, "MRT_EXT_DISPERSION needs MRT_EXT_REFRACTION");
    NEED(!(flags & MRT_EXT_CONTINUOUS_SPECTRUM) || (flags & MRT_EXT_DISPERSION), MRT_E_INVALID, "MRT_EXT_CONTINUOUS_SPECTRUM needs MRT_EXT_DISPERSION");
    NEED(std::isfinite(dispersion) && std::fabs(dispersion) <= 0.5f, MRT_E_INVALID, "dispersion %g outside [-0.5, 0.5]", (double)dispersion);
    c->S.extFlags = flags; c->S.dispersion = (flags & MRT_EXT_DISPERSION) ? dispersion : 0.0f;
    return MRT_OK;
}
int mrt_build_alias_table(const float* pdf, uint32_t count, float* prob, uint32_t* alias) {
    if (!pdf || !prob || !alias) return MRT_E_INVALID;
    build_alias_table(pdf, count, prob, alias);
    return MRT_OK;
}
int mrt_set_pipeline(mrt_ctx* c, uint32_t pipeline) {
    NEED(c && pipeline <= MRT_PIPELINE_STAGED, MRT_E_INVALID, "bad pipeline");
    c->cfg.pipeline = pipeline;
    return MRT_OK;
}

}  // extern "C"

// ---- stage entry points, frame rendering, read-back -------------------------------------------------------------
#define NEED_READY(c) do { NEED(c, MRT_E_INVALID, "ctx is NULL"); NEED((c)->haveScene, MRT_E_STATE, "no scene: call mrt_set_scene first"); \
    NEED((c)->dRays, MRT_E_STATE, "no size: call mrt_resize first"); CK(cudaSetDevice((c)->cfg.device)); } while (0)
#define LAUNCHED(c, n) do { (c)->stats.kernel_launches += (n); CK(cudaGetLastError()); } while (0)

static inline const float4* slot_tile(const mrt_ctx* c, uint32_t slot) { return c->dNoisePool + (slot % MRT_NOISE_SLOTS) * kTileFloat4; }


// ---- the wavefront batch: work items (frame, pixel) handed to path slots dynamically, finished samples parked in a
// per-sample buffer and folded into the image in frame order at the end.  The batch is cut into sub-batches so that the
// sample buffer (16 B per sample) stays within a quarter of the free device memory.
//
// How the waves are issued:
//   * default — ONE CUDA graph per sub-batch: first shade -> WHILE { trace, shade, trace, shade } -> resolve.  Block 0 of
//     the trace kernel ends the loop (cudaGraphSetConditional) when nothing is queued and every sample of the batch has
//     been parked, so no empty wave is launched and the host issues a single cudaGraphLaunch.  The instantiated graph is
//     kept and replayed as long as the kernel arguments do not change.
//   * profile mode (mrt_set_profile: a CUDA-event pair around every trace launch), MRT_WF_GRAPH=0 and the experiments
//     build's slot groups — a host loop that runs at most 2 x 8 waves ahead of the device and stops when the trace kernel
//     has written the batch's serial number to a mapped host word.
// A loop that reaches the proven bound frames x L + L + 2 without finishing ends anyway; k_wf_resolve_samples then
// leaves the failure word that check_drained reports.
static int env_knob(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
#ifdef MRT_EXPERIMENTS
#define EXP_KNOB(name, dflt) env_knob(name, dflt)
#else
#define EXP_KNOB(name, dflt) (dflt)
#endif

static int build_wave_graph(mrt_ctx* c, const RenderParams& q, WfBuffers b, bool count, const std::string& key) {
    if (c->wfGraph) { cudaGraphExecDestroy(c->wfGraph); c->wfGraph = nullptr; c->wfGraphKey.clear(); }
    cudaGraph_t g = nullptr, body = nullptr, tmp = nullptr;
    cudaGraphConditionalHandle handle = 0;
    cudaError_t e = cudaGraphCreate(&g, 0);
    bool capMain = false, capBody = false;
    if (e == cudaSuccess) e = cudaGraphConditionalHandleCreate(&handle, g, 1, cudaGraphCondAssignDefault);
    b.condHandle = (unsigned long long)handle; b.hostStop = nullptr; b.batchSerial = 0;
    if (e == cudaSuccess) { e = cudaStreamBeginCaptureToGraph(c->stream, g, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed); capMain = e == cudaSuccess; }
    if (e == cudaSuccess) {
        launch_wf_shade(c->stream, q, b, 0, true, count);
        launch_wf_sort(c->stream, b, 0);
        e = cudaGetLastError();
    }
    cudaGraphNode_t loop = nullptr;
    if (e == cudaSuccess) {
        cudaStreamCaptureStatus st; const cudaGraphNode_t* deps = nullptr; size_t nDeps = 0;
        e = cudaStreamGetCaptureInfo_v2(c->stream, &st, nullptr, nullptr, &deps, &nDeps);
        cudaGraphNodeParams np = { cudaGraphNodeTypeConditional };
        np.conditional.handle = handle; np.conditional.type = cudaGraphCondTypeWhile; np.conditional.size = 1;
        if (e == cudaSuccess) e = cudaGraphAddNode(&loop, g, deps, nDeps, &np);
        if (e == cudaSuccess) { body = np.conditional.phGraph_out[0]; e = cudaStreamUpdateCaptureDependencies(c->stream, &loop, 1, cudaStreamSetCaptureDependencies); }
    }
    if (e == cudaSuccess) { e = cudaStreamBeginCaptureToGraph(c->captureStream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed); capBody = e == cudaSuccess; }
    if (e == cudaSuccess) {
        for (uint32_t w = 0; w < 2u; ++w) {
            launch_wf_trace(c->captureStream, q, b, w, c->countVisits);
            launch_wf_shade(c->captureStream, q, b, w, false, count);
            launch_wf_sort(c->captureStream, b, (w + 1u) & 1u);
        }
        e = cudaGetLastError();
    }
    if (capBody) { cudaError_t e2 = cudaStreamEndCapture(c->captureStream, &tmp); if (e == cudaSuccess) e = e2; }
    if (e == cudaSuccess) { launch_wf_resolve_samples(c->stream, q, b); e = cudaGetLastError(); }
    if (capMain) { cudaError_t e2 = cudaStreamEndCapture(c->stream, &tmp); if (e == cudaSuccess) e = e2; }
    if (e == cudaSuccess) e = cudaGraphInstantiate(&c->wfGraph, g, 0);
    if (g) cudaGraphDestroy(g);
    if (e != cudaSuccess) {
        c->wfGraph = nullptr;
        cudaGetLastError();
        return fail(MRT_E_CUDA, "building the wave graph failed: %s", cudaGetErrorString(e));
    }
    c->wfGraphKey = key;
    return MRT_OK;
}

static int render_wavefront(mrt_ctx* c, const RenderParams& p, uint32_t n_frames) {
    WfBuffers b = c->wf;
    b.waveLog = nullptr;
    b.pixels = ((c->W + 7u) / 8u) * ((p.ownedRows + 3u) / 4u) * 32u;
    b.slots = b.pixels * c->wfSlotMult;
    const bool count = c->cfg.count_rays != 0;
    size_t freeB = 0, totalB = 0;
    CK(cudaMemGetInfo(&freeB, &totalB));
    const size_t perFrame = (size_t)std::max(1u, b.pixels) * sizeof(float4);
    const size_t budget = std::max(perFrame, std::min<size_t>((freeB + c->sampleCap) / 4, (size_t)24 << 30));
    const uint32_t maxSub = (uint32_t)std::max<size_t>(1, std::min<size_t>(budget / perFrame, n_frames));
    if (c->sampleCap < (size_t)maxSub * perFrame) {
        CK(cudaStreamSynchronize(c->stream));
        cudaFree(c->dSamples); c->dSamples = nullptr; c->sampleCap = 0;
        CK(cudaMalloc(&c->dSamples, (size_t)maxSub * perFrame));
        c->sampleCap = (size_t)maxSub * perFrame;
    }
    b.samples = c->dSamples;
    c->profUsed = 0;
    // ray ordering (experiments build: MRT_WF_SORT = grid bits per axis, 0 = off; MRT_WF_SORT_MODE bit 0 = octant-major keys).
    // Measured in pass r2-A (profiles/r2_a_*): -3 % trace time at 19.05 instead of 18.68 lanes per instruction, against
    // +0.3 ms per wave for counting and scattering the keys — off, and compiled out of the production library.
    static const int sortBits = std::min(3, std::max(0, EXP_KNOB("MRT_WF_SORT", 0)));
    static const int sortMode = EXP_KNOB("MRT_WF_SORT_MODE", 0);
    b.sortBits = (uint32_t)sortBits; b.sortMode = (uint32_t)sortMode;
    for (int a = 0; a < 3; ++a) {
        const float ext = std::max(c->sceneHi[a] - c->sceneLo[a], 1e-20f);
        b.sortLo[a] = c->sceneLo[a]; b.sortScale[a] = float(1u << b.sortBits) / ext;
    }
    // waves per sub-batch of nf frames: greedy assignment finishes within (total segments / slots) + L waves, plus
    // one wave per padding pixel a slot may draw
    const uint32_t Lp = c->cfg.max_path_length;
    // MRT_WF_MAX_WAVES (tests only): cut the schedule short to exercise the failure flag of check_drained
    static const long long waveCap = getenv("MRT_WF_MAX_WAVES") ? atoll(getenv("MRT_WF_MAX_WAVES")) : 0;
    auto waves_for = [Lp](uint32_t nf) {
        const uint64_t bound = (uint64_t)nf * Lp + Lp + 2;
        return waveCap > 0 ? std::min<uint64_t>(bound, (uint64_t)waveCap) : bound;
    };
    uint64_t wavesTotal = 0;
    for (uint32_t f0 = 0; f0 < n_frames; f0 += maxSub) wavesTotal += 4 * waves_for(std::min(maxSub, n_frames - f0));   // x4: up to 4 slot groups
    if (c->profile) {
        while (c->profEvents.size() < 2 * wavesTotal) { cudaEvent_t e; CK(cudaEventCreate(&e)); c->profEvents.push_back(e); }
        if (c->waveCountCap < wavesTotal) {
            cudaFree(c->dWaveCounts); c->dWaveCounts = nullptr; c->waveCountCap = 0;
            CK(cudaMalloc(&c->dWaveCounts, wavesTotal * 2 * sizeof(unsigned int)));
            c->waveCountCap = wavesTotal;
        }
    }
    // Slot groups (experiments build, MRT_WF_SPLIT=2..4): the slots are cut into equal parts, each with its own queues and
    // stream, so that one group's kernels could fill the SMs freed while another group's persistent trace kernel drains.
    // Measured on B200 (profiles/r1_k_sweep.log): 2 groups = no change, 3-4 groups 4-8 % slower; with the 64-register
    // shade kernel 2 groups gain 1-2.6 %.  The production path runs one group on one stream.
    static const int splitEnv = EXP_KNOB("MRT_WF_SPLIT", 1);
    uint32_t groups = (uint32_t)std::min(4, std::max(1, splitEnv));
    while (groups > 1 && b.slots / groups < 262144u) --groups;                     // small images: not worth splitting
    uint32_t per = ((b.slots + groups - 1u) / groups + 255u) / 256u * 256u;
    static const int graphEnv = env_knob("MRT_WF_GRAPH", 1);
    const bool useGraph = graphEnv != 0 && c->waveSchedule == MRT_SCHEDULE_GRAPH && !c->profile && groups == 1;
    const uint32_t perWave = 2u + (b.sortBits ? 1u : 0u);                            // kernels per wave: trace, shade (, sort)
    uint64_t launched = 0, pw = 0;
    c->kernelsPerWave = perWave;
    for (uint32_t f0 = 0; f0 < n_frames; f0 += maxSub) {
        const uint32_t nf = std::min(maxSub, n_frames - f0);
        RenderParams q = p;
        q.frames = c->dFrameTable + f0; q.nFrames = nf;
        {   // MRT_WF_REGION (experiments build): 32 = per-tile counters claimed per warp (default), 256 = per shade block, 0 = one global counter
            static const int rsEnv = EXP_KNOB("MRT_WF_REGION", 32);
            b.regionSize = rsEnv == 32 ? 32u : (rsEnv <= 0 ? std::max(1u, b.pixels) : (uint32_t)((rsEnv + 255) / 256 * 256));
            b.regions = (b.pixels + b.regionSize - 1u) / b.regionSize;
        }
        const uint64_t waves = waves_for(nf);
        b.expected = c->W * p.ownedRows * nf;
        b.waveLimit = (uint32_t)std::min<uint64_t>(waves, 0xFFFFFFFFull);
        b.condHandle = 0; b.hostStop = nullptr; b.batchSerial = 0;
        b.allCounts = b.counts; b.groups = groups; b.slotBase = 0;
        // per group: counts[4], heads[2]; word 7 of group 0 is the failure word (check_drained reports and clears it)
        CK(cudaMemsetAsync(b.counts, 0, 7 * sizeof(unsigned int), c->stream));
        CK(cudaMemsetAsync(b.counts + 8, 0, 24 * sizeof(unsigned int), c->stream));
        CK(cudaMemsetAsync(b.ctl, 0, MRT_CTL_BATCH_WORDS * sizeof(unsigned int), c->stream));
        CK(cudaMemsetAsync(b.regionNext, 0, (size_t)b.regions * sizeof(unsigned int), c->stream));
        if (b.sortBits) CK(cudaMemsetAsync(b.sortHist, 0, 4 * (size_t)wf_sort_buckets(b.sortBits) * sizeof(unsigned int), c->stream));
        if (useGraph) {
            std::string key(sizeof q + sizeof b + 2, '\0');
            std::memcpy(&key[0], &q, sizeof q); std::memcpy(&key[sizeof q], &b, sizeof b);
            key[sizeof q + sizeof b] = count ? 1 : 0; key[sizeof q + sizeof b + 1] = c->countVisits ? 1 : 0;
            if (!c->wfGraph || key != c->wfGraphKey) { int rc = build_wave_graph(c, q, b, count, key); if (rc != MRT_OK) return rc; }
            CK(cudaGraphLaunch(c->wfGraph, c->stream));
            launched += 2 + (b.sortBits ? 1 : 0);                                 // first shade (+ sort), resolve; the waves are counted on the device
            continue;
        }
        WfBuffers gb[4];
        for (uint32_t g = 0; g < groups; ++g) {
            const uint32_t at = std::min(b.slots, g * per), n = std::min(per, b.slots - at);
            WfBuffers h = b;
            h.P0 += at; h.P1 += at; h.P2 += at; h.P3 += at; h.H += at; h.S0 += at; h.S1 += at; h.visible += at; h.item += at;
            for (int k = 0; k < 2; ++k) { h.queueExt[k] += at; h.queueShadow[k] += at; h.sortSlot[k] += at; h.sortKey[k] += at; }
            h.sortHist = b.sortHist + (size_t)g * 4 * wf_sort_buckets(b.sortBits);
            h.counts = b.counts + 8 * g; h.heads = h.counts + 4;
            h.slots = n; h.slotBase = at;
            gb[g] = h;
        }
        const bool poll = groups == 1;
        const uint32_t serial = ++c->batchSerial;
        if (poll) { gb[0].hostStop = c->hStop; gb[0].batchSerial = serial; }
        auto stream_of = [&](uint32_t g) { return (c->profile || groups == 1) ? c->stream : c->groupStream[g]; };
        const bool forked = stream_of(0) != c->stream;
        if (forked) {
            CK(cudaEventRecord(c->evFork, c->stream));
            for (uint32_t g = 0; g < groups; ++g) CK(cudaStreamWaitEvent(c->groupStream[g], c->evFork, 0));
        }
        for (uint32_t g = 0; g < groups; ++g) { launch_wf_shade(stream_of(g), q, gb[g], 0, true, count); launch_wf_sort(stream_of(g), gb[g], 0); }
        launched += groups * (perWave - 1u);
        const uint32_t K = 8;                                                     // poll every K waves, run at most 2K waves ahead
        for (uint64_t w = 0; w < waves; ++w) {
            if (poll && w % K == 0) {
                if (w >= 2 * K) {
                    CK(cudaEventSynchronize(c->pollEvents[(w / K - 2) % 4]));
                    if (*c->hStop == serial) break;
                }
                CK(cudaEventRecord(c->pollEvents[(w / K) % 4], c->stream));
            }
            for (uint32_t g = 0; g < groups; ++g, ++pw) {
                cudaStream_t st = stream_of(g);
                gb[g].waveLog = nullptr;
                if (c->profile) {
                    gb[g].waveLog = c->dWaveCounts + 2 * pw;        // the trace kernel notes its queue sizes there
                    CK(cudaEventRecord(c->profEvents[2 * pw], st));
                }
                launch_wf_trace(st, q, gb[g], (uint32_t)(w & 1u), c->countVisits);
                if (c->profile) CK(cudaEventRecord(c->profEvents[2 * pw + 1], st));
                launch_wf_shade(st, q, gb[g], (uint32_t)w, false, count);
                launch_wf_sort(st, gb[g], (uint32_t)((w + 1u) & 1u));
                launched += perWave;
            }
        }
        if (forked) {
            for (uint32_t g = 0; g < groups; ++g) {
                CK(cudaEventRecord(c->evJoin[g], c->groupStream[g]));
                CK(cudaStreamWaitEvent(c->stream, c->evJoin[g], 0));
            }
        }
        launch_wf_resolve_samples(c->stream, q, b);
        launched += 1;
    }
    if (c->profile) c->profUsed = 2 * pw;
    c->stats.kernel_launches += launched;
    CK(cudaGetLastError());
    return MRT_OK;
}

extern "C" {

int mrt_raygen(mrt_ctx* c, uint32_t frame_index) {
    NEED_READY(c);
    NvtxRange range("Ray generator");
    launch_raygen(c->stream, c->dRays, slot_tile(c, frame_index), c->grid());               // Renderer.mm:508-512
    LAUNCHED(c, 1);
    return MRT_OK;
}

int mrt_intersect(mrt_ctx* c, uint32_t which) {
    NEED_READY(c);
    NEED(which == MRT_BUF_RAYS || which == MRT_BUF_LIGHT_RAYS, MRT_E_INVALID, "mrt_intersect: which must be MRT_BUF_RAYS or MRT_BUF_LIGHT_RAYS");
    const void* rays = which == MRT_BUF_RAYS ? (const void*)c->dRays : (const void*)c->dLightRays;
    uint32_t stride = which == MRT_BUF_RAYS ? sizeof(mrt_ray) : sizeof(mrt_light_sampling_ray);  // setRayStride, Renderer.mm:519,545
    unsigned long long* live = c->cfg.count_rays ? c->dCounters + (which == MRT_BUF_RAYS ? 0 : 1) : nullptr;
    NvtxRange range(which == MRT_BUF_RAYS ? "Intersect rays" : "Intersect light-sampling rays");
    CK(cudaMemsetAsync(c->dHead, 0, sizeof(unsigned int), c->stream));
    launch_intersect_persistent(c->stream, c->S, rays, stride, c->W * c->H, c->dIsects, c->grid(), true, live, c->dHead);
    LAUNCHED(c, 1);
    return MRT_OK;
}

int mrt_shade(mrt_ctx* c, uint32_t frame_index, uint32_t bounce_iteration) {
    NEED_READY(c);
    NvtxRange range("Intersection handler");
    launch_shade(c->stream, c->S, c->dIsects, c->dRays, c->dLightRays, slot_tile(c, frame_index + bounce_iteration),  // Renderer.mm:538
                 frame_index, c->cfg.max_path_length, c->grid());
    LAUNCHED(c, 1);
    return MRT_OK;
}

int mrt_shadow_resolve(mrt_ctx* c) {
    NEED_READY(c);
    NvtxRange range("Light sampling");
    launch_shadow_resolve(c->stream, c->dIsects, c->dRays, c->dLightRays, c->grid());
    LAUNCHED(c, 1);
    return MRT_OK;
}

int mrt_accumulate(mrt_ctx* c, uint32_t frame_index) {
    NEED_READY(c);
    NvtxRange range("Final image accumulation");
    launch_accumulate(c->stream, c->dImage, c->dRays, frame_index, c->cfg.accumulate_mode, c->grid());
    LAUNCHED(c, 1);
    c->nextFrame = frame_index + 1;
    return MRT_OK;
}

// Renderer.mm:504-584: 1 + 4*L + 1 dispatches
int mrt_render_frame_staged(mrt_ctx* c, uint32_t frame_index) {
    int rc = mrt_raygen(c, frame_index);
    for (uint32_t i = 0; rc == MRT_OK && i < c->cfg.max_path_length; ++i) {
        if ((rc = mrt_intersect(c, MRT_BUF_RAYS))) break;
        if ((rc = mrt_shade(c, frame_index, i))) break;
        if ((rc = mrt_intersect(c, MRT_BUF_LIGHT_RAYS))) break;
        rc = mrt_shadow_resolve(c);
    }
    if (rc == MRT_OK) rc = mrt_accumulate(c, frame_index);
    if (rc == MRT_OK) c->stats.frames_rendered += 1;
    return rc;
}

int mrt_render_frames(mrt_ctx* c, uint32_t first_frame, uint32_t n_frames, uint32_t frame_stride, const float* noise_tiles) {
    NEED_READY(c);
    if (n_frames == 0) return MRT_OK;
    NvtxRange range("performRaytracing (batch)");
    if (frame_stride == 0) frame_stride = 1;
    NEED((uint64_t)first_frame + (uint64_t)(n_frames - 1) * frame_stride < 0xFFFFFFFFull, MRT_E_INVALID, "frame index overflow");
    const uint32_t span = (n_frames - 1) * frame_stride + 1;              // tiles supplied: one per frame first..last
    const bool refresh = noise_tiles != nullptr && c->cfg.animate_noise;
    CK(cudaEventRecord(c->ev0, c->stream));

    if (c->cfg.pipeline == MRT_PIPELINE_STAGED) {
        // the reference's own shape: per frame refresh slot f%3 (Renderer.mm:492-496), then 2 + 4*L dispatches
        for (uint32_t k = 0; k < span; ++k) {
            uint32_t f = first_frame + k;
            if (refresh) CK(cudaMemcpyAsync(c->dNoisePool + (f % 3u) * kTileFloat4, noise_tiles + (size_t)k * MRT_NOISE_TILE_FLOATS, kTileBytes,
                                            cudaMemcpyDefault, c->stream));
            if (k % frame_stride == 0) { int rc = mrt_render_frame_staged(c, f); if (rc) return rc; }
        }
    } else {
        int rc = ensure_pool(c, MRT_NOISE_SLOTS + (refresh ? span : 0));
        if (rc) return rc;
        rc = ensure_table(c, n_frames);
        if (rc) return rc;
        if (refresh) CK(cudaMemcpyAsync(c->dNoisePool + MRT_NOISE_SLOTS * kTileFloat4, noise_tiles, (size_t)span * kTileBytes,
                                        cudaMemcpyDefault, c->stream));
        CK(cudaEventSynchronize(c->tableCopied));                         // previous batch's table has left the pinned buffer
        uint32_t slotTile[3] = {0, 1, 2};
        uint32_t e = 0;
        for (uint32_t k = 0; k < span; ++k) {
            uint32_t f = first_frame + k;
            if (refresh) slotTile[f % 3u] = MRT_NOISE_SLOTS + k;
            if (k % frame_stride == 0) { FrameEntry& fe = c->hFrameTable[e++]; fe.frame = f; fe.tile[0] = slotTile[0]; fe.tile[1] = slotTile[1]; fe.tile[2] = slotTile[2]; }
        }
        CK(cudaMemcpyAsync(c->dFrameTable, c->hFrameTable, n_frames * sizeof(FrameEntry), cudaMemcpyHostToDevice, c->stream));
        CK(cudaEventRecord(c->tableCopied, c->stream));
        RenderParams p;
        std::memset(&p, 0, sizeof p);                                     // padding too: the wave graph is keyed on these bytes
        p.S = c->S; p.noisePool = c->dNoisePool; p.frames = c->dFrameTable; p.nFrames = n_frames; p.image = c->dImage;
        p.g = c->grid(); p.ownedRows = c->ownedRows(); p.maxPathLength = c->cfg.max_path_length; p.accumMode = c->cfg.accumulate_mode;
        p.counters = c->cfg.count_rays ? c->dCounters : nullptr;
        CK(cudaEventRecord(c->evK0, c->stream));
        if (c->cfg.pipeline == MRT_PIPELINE_MEGAKERNEL) {
            NEED(c->S.extFlags == 0, MRT_E_STATE, "the optics extension (mrt_set_extensions) runs on the wavefront and staged pipelines only");
            p.counters = c->dCounters;
            launch_render_fused(c->stream, p, c->cfg.count_rays != 0);
            LAUNCHED(c, 1);
        } else {
            // queue-compacted wavefront (the production pipeline)
            int rcw = render_wavefront(c, p, n_frames);
            if (rcw != MRT_OK) return rcw;
        }
        CK(cudaEventRecord(c->evK1, c->stream));
        c->kernelTimed = true;
        // leave the three slots as the reference's would be after the last frame
        for (uint32_t s = 0; s < 3 && refresh; ++s)
            if (slotTile[s] != s) CK(cudaMemcpyAsync(c->dNoisePool + s * kTileFloat4, c->dNoisePool + slotTile[s] * kTileFloat4, kTileBytes,
                                                     cudaMemcpyDeviceToDevice, c->stream));
        c->stats.frames_rendered += n_frames;
        c->nextFrame = first_frame + span;
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    return MRT_OK;
}

// Failure detection of the wavefront schedule: k_wf_resolve_samples leaves a non-zero word when a batch ended with rays
// still queued or work items unclaimed (the wave bound of mrt_render_frames would have to be wrong).  Every call that
// hands results to the caller checks it after its stream synchronisation, so a short render can never pass silently.
static int check_drained(mrt_ctx* c) {
    if (!c->wf.counts) return MRT_OK;
    unsigned int undrained = 0;
    CK(cudaMemcpy(&undrained, c->wf.counts + 7, sizeof undrained, cudaMemcpyDeviceToHost));
    if (undrained) {            // reported once: the word is cleared so that later, complete renders are not failed for it
        CK(cudaMemset(c->wf.counts + 7, 0, sizeof undrained));
        return fail(MRT_E_STATE, "wavefront batch ended with %u rays still queued (wave bound too small)", undrained);
    }
    return MRT_OK;
}

int mrt_sync(mrt_ctx* c) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    return check_drained(c);
}

int mrt_event_record(mrt_ctx* c, uint32_t slot) {
    NEED(c && slot < 8, MRT_E_INVALID, "event slot out of range");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaEventRecord(c->user[slot], c->stream));
    return MRT_OK;
}
int mrt_event_elapsed_ms(mrt_ctx* c, uint32_t a, uint32_t b, float* ms) {
    NEED(c && ms && a < 8 && b < 8, MRT_E_INVALID, "bad argument");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaEventSynchronize(c->user[b]));
    CK(cudaEventElapsedTime(ms, c->user[a], c->user[b]));
    return MRT_OK;
}

int mrt_read_image(mrt_ctx* c, float* rgba) {
    NEED(c && rgba, MRT_E_INVALID, "NULL argument");
    NEED(c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(rgba, c->dImage, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return check_drained(c);
}
int mrt_scale_image(mrt_ctx* c, float scale) {
    NEED(c && c->dImage, MRT_E_STATE, "no image");
    CK(cudaSetDevice(c->cfg.device));
    launch_scale_image(c->stream, c->dImage, scale, c->W * c->H);
    LAUNCHED(c, 1);
    return MRT_OK;
}
void mrt_display_params_default(mrt_display_params* p) {
    if (!p) return;
    p->tone_mapping = 0; p->manual_srgb = 0;                       // Raytracing.h:11-12
    p->comparison_mode = MRT_COMPARE_DISABLED; p->comparison_scale = MRT_COMPARISON_SCALE;   // Raytracing.h:32-33
}

int mrt_display(mrt_ctx* c, const mrt_display_params* params, const float* reference_rgba, uint32_t ref_width, uint32_t ref_height,
                float* out_rgba) {
    NEED(c && params && out_rgba, MRT_E_INVALID, "NULL argument");
    NEED(c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    NEED(params->comparison_mode <= MRT_COMPARE_LUMINANCE, MRT_E_INVALID, "comparison mode %u out of range", params->comparison_mode);
    NEED(reference_rgba == nullptr || (ref_width >= 1 && ref_height >= 1), MRT_E_INVALID, "reference image without a size");
    CK(cudaSetDevice(c->cfg.device));
    const size_t n = (size_t)c->W * c->H;
    float4 *dOut = nullptr, *dRef = nullptr;
    CK(cudaMalloc(&dOut, n * sizeof(float4)));
    cudaError_t e = cudaSuccess;
    if (reference_rgba) {
        const size_t rb = (size_t)ref_width * ref_height * sizeof(float4);
        e = cudaMalloc(&dRef, rb);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dRef, reference_rgba, rb, cudaMemcpyDefault, c->stream);
    }
    if (e == cudaSuccess) {
        launch_display(c->stream, c->dImage, dRef, dOut, c->W, c->H, ref_width, ref_height, params->tone_mapping, params->manual_srgb,
                       params->comparison_mode, params->comparison_scale);
        LAUNCHED(c, 1);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_rgba, dOut, n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(dOut); cudaFree(dRef);
    CK(e);
    return MRT_OK;
}

// ---- checkpoint: "MRTCKPT1", W, H, nextFrame, accumulate mode, image, the three noise slots ------------------------
static const char kCkptMagic[8] = {'M', 'R', 'T', 'C', 'K', 'P', 'T', '1'};

int mrt_checkpoint_save(mrt_ctx* c, const char* path) {
    NEED(c && path, MRT_E_INVALID, "NULL argument");
    NEED(c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    CK(cudaSetDevice(c->cfg.device));
    const size_t n = (size_t)c->W * c->H;
    std::vector<float> image(4 * n), noise((size_t)MRT_NOISE_SLOTS * MRT_NOISE_TILE_FLOATS);
    CK(cudaMemcpyAsync(image.data(), c->dImage, n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(noise.data(), c->dNoisePool, noise.size() * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    FILE* f = fopen(path, "wb");
    NEED(f, MRT_E_IO, "cannot open %s for writing", path);
    const uint32_t head[4] = {c->W, c->H, c->nextFrame, c->cfg.accumulate_mode};
    bool ok = fwrite(kCkptMagic, 1, 8, f) == 8 && fwrite(head, sizeof head, 1, f) == 1 &&
              fwrite(image.data(), sizeof(float), image.size(), f) == image.size() &&
              fwrite(noise.data(), sizeof(float), noise.size(), f) == noise.size();
    ok = (fclose(f) == 0) && ok;
    NEED(ok, MRT_E_IO, "short write to %s", path);
    return MRT_OK;
}

int mrt_checkpoint_load(mrt_ctx* c, const char* path, uint32_t* frames_accumulated) {
    NEED(c && path, MRT_E_INVALID, "NULL argument");
    FILE* f = fopen(path, "rb");
    NEED(f, MRT_E_IO, "cannot open %s", path);
    char magic[8]; uint32_t head[4] = {0, 0, 0, 0};
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, kCkptMagic, 8) == 0 && fread(head, sizeof head, 1, f) == 1;
    ok = ok && head[0] >= 2 && head[1] >= 2 && (uint64_t)head[0] * head[1] < (1ull << 31) && head[3] <= MRT_ACCUM_REPLACE;
    std::vector<float> image, noise((size_t)MRT_NOISE_SLOTS * MRT_NOISE_TILE_FLOATS);
    if (ok) {
        image.resize((size_t)4 * head[0] * head[1]);
        ok = fread(image.data(), sizeof(float), image.size(), f) == image.size() &&
             fread(noise.data(), sizeof(float), noise.size(), f) == noise.size();
    }
    fclose(f);
    NEED(ok, MRT_E_IO, "%s is not a complete checkpoint", path);
    if (c->W != head[0] || c->H != head[1] || !c->dImage) { int rc = mrt_resize(c, head[0], head[1]); if (rc != MRT_OK) return rc; }
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(c->dImage, image.data(), image.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->dNoisePool, noise.data(), noise.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->nextFrame = head[2]; c->cfg.accumulate_mode = head[3];
    if (frames_accumulated) *frames_accumulated = head[2];
    return MRT_OK;
}

int mrt_clear_image(mrt_ctx* c) {
    NEED(c && c->dImage, MRT_E_STATE, "no image");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemsetAsync(c->dImage, 0, (size_t)c->W * c->H * sizeof(float4), c->stream));
    return MRT_OK;
}
int mrt_image_device_ptr(mrt_ctx* c, void** device_ptr, size_t* bytes) {
    NEED(c && device_ptr && c->dImage, MRT_E_STATE, "no image");
    *device_ptr = c->dImage;
    if (bytes) *bytes = (size_t)c->W * c->H * sizeof(float4);
    return MRT_OK;
}

// ---- multi-GPU plumbing: everything below is enqueued on the ctx stream, so a collective issued on mrt_stream_handle()
// between two of these calls needs no host synchronisation ------------------------------------------------------------
int mrt_stream_handle(mrt_ctx* c, void** cuda_stream) {
    NEED(c && cuda_stream, MRT_E_INVALID, "NULL argument");
    *cuda_stream = (void*)c->stream;
    return MRT_OK;
}
int mrt_scale_buffer(mrt_ctx* c, void* device_rgba, size_t pixels, float scale) {
    NEED(c && device_rgba, MRT_E_INVALID, "NULL argument");
    NEED(pixels < (1ull << 32), MRT_E_INVALID, "too many pixels");
    CK(cudaSetDevice(c->cfg.device));
    if (pixels) { launch_scale_image(c->stream, static_cast<float4*>(device_rgba), scale, (uint32_t)pixels); LAUNCHED(c, 1); }
    return MRT_OK;
}
int mrt_shard_layout(mrt_ctx* c, uint32_t shards, uint32_t* rows_max, size_t* bytes_per_shard) {
    NEED(c && c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    NEED(shards >= 1, MRT_E_INVALID, "shards must be >= 1");
    const uint32_t rm = (c->H + shards - 1) / shards;
    if (rows_max) *rows_max = rm;
    if (bytes_per_shard) *bytes_per_shard = (size_t)rm * c->W * sizeof(float4);
    return MRT_OK;
}
int mrt_pack_rows(mrt_ctx* c, void* device_dst) {
    NEED(c && device_dst, MRT_E_INVALID, "NULL argument");
    NEED(c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    CK(cudaSetDevice(c->cfg.device));
    const size_t rowBytes = (size_t)c->W * sizeof(float4);
    const uint32_t rows = c->ownedRows();
    if (rows) CK(cudaMemcpy2DAsync(device_dst, rowBytes, c->dImage + (size_t)c->rowFirst * c->W, rowBytes * c->rowStride, rowBytes, rows,
                                   cudaMemcpyDeviceToDevice, c->stream));
    return MRT_OK;
}
int mrt_unpack_rows(mrt_ctx* c, const void* device_src, uint32_t shards) {
    NEED(c && device_src, MRT_E_INVALID, "NULL argument");
    NEED(c->dImage, MRT_E_STATE, "no size: call mrt_resize first");
    NEED(shards >= 1, MRT_E_INVALID, "shards must be >= 1");
    CK(cudaSetDevice(c->cfg.device));
    const size_t rowBytes = (size_t)c->W * sizeof(float4);
    const uint32_t rowsMax = (c->H + shards - 1) / shards;
    for (uint32_t g = 0; g < shards && g < c->H; ++g) {
        const uint32_t rows = (c->H - g + shards - 1) / shards;
        CK(cudaMemcpy2DAsync(c->dImage + (size_t)g * c->W, rowBytes * shards, static_cast<const char*>(device_src) + (size_t)g * rowsMax * rowBytes,
                             rowBytes, rowBytes, rows, cudaMemcpyDeviceToDevice, c->stream));
    }
    return MRT_OK;
}

static int buffer_of(mrt_ctx* c, uint32_t which, void** p, size_t* bytes) {
    size_t n = (size_t)c->W * c->H;
    switch (which) {
        case MRT_BUF_RAYS: *p = c->dRays; *bytes = n * sizeof(mrt_ray); return MRT_OK;
        case MRT_BUF_LIGHT_RAYS: *p = c->dLightRays; *bytes = n * sizeof(mrt_light_sampling_ray); return MRT_OK;
        case MRT_BUF_INTERSECTIONS: *p = c->dIsects; *bytes = n * sizeof(mrt_intersection); return MRT_OK;
        case MRT_BUF_IMAGE: *p = c->dImage; *bytes = n * sizeof(float4); return MRT_OK;
    }
    return fail(MRT_E_INVALID, "unknown buffer %u", which);
}
int mrt_read_buffer(mrt_ctx* c, uint32_t which, void* dst, size_t bytes) {
    NEED(c && dst && c->dRays, MRT_E_STATE, "no buffers: call mrt_resize first");
    void* p = nullptr; size_t cap = 0; int rc = buffer_of(c, which, &p, &cap); if (rc) return rc;
    NEED(bytes <= cap, MRT_E_INVALID, "read of %zu bytes exceeds buffer (%zu)", bytes, cap);
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return MRT_OK;
}
int mrt_write_buffer(mrt_ctx* c, uint32_t which, const void* src, size_t bytes) {
    NEED(c && src && c->dRays, MRT_E_STATE, "no buffers: call mrt_resize first");
    void* p = nullptr; size_t cap = 0; int rc = buffer_of(c, which, &p, &cap); if (rc) return rc;
    NEED(bytes <= cap, MRT_E_INVALID, "write of %zu bytes exceeds buffer (%zu)", bytes, cap);
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return MRT_OK;
}

int mrt_get_stats(mrt_ctx* c, mrt_stats* out) {
    NEED(c && out, MRT_E_INVALID, "NULL argument");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    unsigned long long h[4];
    CK(cudaMemcpy(h, c->dCounters, sizeof h, cudaMemcpyDeviceToHost));
    { int rc = check_drained(c); if (rc != MRT_OK) return rc; }
    c->stats.primary_rays = h[0]; c->stats.shadow_rays = h[1]; c->stats.nodes_visited = h[2]; c->stats.triangles_tested = h[3];
    if (c->timed) { float ms = 0; CK(cudaEventElapsedTime(&ms, c->ev0, c->ev1)); c->stats.last_render_ms = ms; }
    if (c->kernelTimed) { float ms = 0; CK(cudaEventElapsedTime(&ms, c->evK0, c->evK1)); c->stats.last_kernel_ms = ms; }
    {   // waves: graph replays count their launches on the device
        unsigned int ctl[MRT_CTL_WORDS] = {};
        if (c->wf.ctl) CK(cudaMemcpy(ctl, c->wf.ctl, sizeof ctl, cudaMemcpyDeviceToHost));
        c->stats.waves_traced = ctl[MRT_CTL_WAVES];
        c->graphKernels = (uint64_t)ctl[MRT_CTL_GRAPH_TRACES] * c->kernelsPerWave;
    }
    c->stats.tail_waves = 0;
    if (c->profUsed) {
        std::vector<unsigned int> wc(c->profUsed);
        CK(cudaMemcpy(wc.data(), c->dWaveCounts, c->profUsed * sizeof(unsigned int), cudaMemcpyDeviceToHost));
        double sum = 0.0; uint32_t n = 0;
        const char* logPath = getenv("MRT_WAVE_LOG");                      // diagnostic: per-wave queue sizes and trace time
        FILE* log = logPath ? fopen(logPath, "w") : nullptr;
        if (log) fprintf(log, "wave,extension_rays,light_rays,trace_ms\n");
        for (size_t w = 0; w < c->profUsed / 2; ++w) {
            if (wc[2 * w] + wc[2 * w + 1] == 0) continue;                 // empty wave: the kernel returned at once
            float ms = 0; CK(cudaEventElapsedTime(&ms, c->profEvents[2 * w], c->profEvents[2 * w + 1]));
            sum += ms; ++n;
            if (2u * wc[2 * w] < c->wf.slots) c->stats.tail_waves += 1;                 // fewer extension rays than half the path slots
            if (log) fprintf(log, "%zu,%u,%u,%.4f\n", w, wc[2 * w], wc[2 * w + 1], ms);
        }
        if (log) fclose(log);
        c->stats.profile_trace_ms = (float)sum; c->stats.profile_trace_launches = n;
    }
    *out = c->stats;
    out->kernel_launches += c->graphKernels;
    return MRT_OK;
}
int mrt_reset_stats(mrt_ctx* c) {
    NEED(c, MRT_E_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemsetAsync(c->dCounters, 0, 4 * sizeof(unsigned long long), c->stream));
    if (c->wf.ctl) CK(cudaMemsetAsync(c->wf.ctl + MRT_CTL_GRAPH_TRACES, 0, 2 * sizeof(unsigned int), c->stream));
    c->stats.kernel_launches = 0; c->stats.frames_rendered = 0; c->graphKernels = 0;
    return MRT_OK;
}

static int intersect_host_impl(mrt_ctx* c, const void* rays, size_t stride, uint32_t count, mrt_intersection* out, bool profile) {
    NEED(c && rays, MRT_E_INVALID, "NULL argument");
    NEED(c->haveScene, MRT_E_STATE, "no scene: call mrt_set_scene first");
    NEED(stride >= 32 && stride % 4 == 0, MRT_E_INVALID, "ray stride must be >= 32 and a multiple of 4");
    if (count == 0) return MRT_OK;
    CK(cudaSetDevice(c->cfg.device));
    void* dR = nullptr; mrt_intersection* dO = nullptr;
    cudaError_t e = cudaMalloc(&dR, (size_t)count * stride);
    if (e == cudaSuccess) e = cudaMalloc(&dO, (size_t)count * sizeof(mrt_intersection));
    if (e == cudaSuccess) e = cudaMemcpyAsync(dR, rays, (size_t)count * stride, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        Grid2D g; g.W = count; g.H = 1; g.rowFirst = 0; g.rowStride = 1;
        launch_intersect(c->stream, c->S, dR, (uint32_t)stride, count, dO, g, false, nullptr, profile ? c->dCounters + 2 : nullptr);
        c->stats.kernel_launches += 1;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && out) e = cudaMemcpyAsync(out, dO, (size_t)count * sizeof(mrt_intersection), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(dR); cudaFree(dO);
    if (e != cudaSuccess) return fail(MRT_E_CUDA, "mrt_intersect_host: %s", cudaGetErrorString(e));
    return MRT_OK;
}
int mrt_intersect_host(mrt_ctx* c, const void* rays, size_t stride_bytes, uint32_t count, mrt_intersection* out) {
    NEED(out, MRT_E_INVALID, "out is NULL");
    return intersect_host_impl(c, rays, stride_bytes, count, out, false);
}
int mrt_profile_traversal(mrt_ctx* c, const void* rays, size_t stride_bytes, uint32_t count) {
    return intersect_host_impl(c, rays, stride_bytes, count, nullptr, true);
}

// ---- caller-side helpers (host only) -------------------------------------------------------------------------
int mrt_scene_create_procedural(const char* name, uint64_t seed, mrt_scene** out) {
    NEED(name && out, MRT_E_INVALID, "NULL argument");
    mrt_scene* s = new (std::nothrow) mrt_scene();
    NEED(s, MRT_E_NOMEM, "out of host memory");
    std::string err;
    if (!make_procedural(name, seed, *s, err)) { delete s; return fail(MRT_E_INVALID, "%s", err.c_str()); }
    *out = s;
    return MRT_OK;
}
int mrt_scene_load_obj(const char* obj_path, mrt_scene** out) {
    NEED(obj_path && out, MRT_E_INVALID, "NULL argument");
    mrt_scene* s = new (std::nothrow) mrt_scene();
    NEED(s, MRT_E_NOMEM, "out of host memory");
    std::string err;
    if (!load_obj(obj_path, *s, err)) { delete s; return fail(MRT_E_IO, "%s", err.c_str()); }
    *out = s;
    return MRT_OK;
}
int mrt_scene_get(const mrt_scene* s, mrt_scene_view* v) {
    NEED(s && v, MRT_E_INVALID, "NULL argument");
    v->vertices = s->vertices.data(); v->vertex_count = (uint32_t)s->vertices.size();
    v->indices = s->indices.data(); v->triangle_count = (uint32_t)s->references.size();
    v->references = s->references.data();
    v->materials = s->materials.data(); v->material_count = (uint32_t)s->materials.size();
    v->lights = s->lights.data(); v->light_count = s->light_count;
    return MRT_OK;
}
void mrt_scene_destroy(mrt_scene* s) { delete s; }

int mrt_bvh_selftest(const mrt_scene* s, uint32_t n_rays, uint64_t seed, mrt_bvh_report* out) {
    NEED(s && out, MRT_E_INVALID, "NULL argument");
    const uint32_t ntri = (uint32_t)s->references.size();
    Bvh8Result bvh;
    build_bvh8(s->vertices.data(), s->indices.data(), ntri, 0, bvh);
    float lo[3] = {-1.0f, -1.0f, -1.0f}, hi[3] = {1.0f, 1.0f, 1.0f};
    if (!s->vertices.empty()) {
        for (int a = 0; a < 3; ++a) { lo[a] = 1e30f; hi[a] = -1e30f; }
        for (const mrt_vertex& v : s->vertices) for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], v.v[a]); hi[a] = std::max(hi[a], v.v[a]); }
    }
    std::mt19937_64 rng(seed);
    auto uni = [&]() { return float(rng() >> 40) * (1.0f / 16777216.0f); };
    std::vector<float> rays(8 * (size_t)n_rays);
    for (uint32_t i = 0; i < n_rays; ++i) {
        float* r = &rays[8 * (size_t)i];
        for (int a = 0; a < 3; ++a) r[a] = lo[a] + (hi[a] - lo[a]) * (uni() * 1.2f - 0.1f);
        float d[3], len;
        do { for (int a = 0; a < 3; ++a) d[a] = uni() * 2.0f - 1.0f; len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); } while (len < 1e-3f || len > 1.0f);
        for (int a = 0; a < 3; ++a) d[a] /= len;
        if (i % 11 == 3) { int a = int(i / 11) % 3; d[a] = 0.0f; }                       // axis-parallel components
        if (i % 13 == 5) { int a = int(i / 13) % 3; d[0] = d[1] = d[2] = 0.0f; d[a] = (i & 1) ? 1.0f : -1.0f; }
        if (i % 17 == 7 && !s->vertices.empty()) {                                      // origin exactly on a vertex plane
            const mrt_vertex& v = s->vertices[rng() % s->vertices.size()]; int a = int(i / 17) % 3; r[a] = v.v[a];
        }
        r[3] = (i % 19 == 2) ? 0.25f : 0.0f;
        r[4] = d[0]; r[5] = d[1]; r[6] = d[2];
        r[7] = (i % 23 == 4) ? 0.8f : INFINITY;
    }
    uint64_t nv = 0, tt = 0;
    out->mismatches = bvh8_selftest(bvh, rays.data(), n_rays, &nv, &tt);
    out->nodes = (uint32_t)bvh.nodes.size(); out->depth = bvh.depth; out->triangles = (uint32_t)bvh.tris.size();
    out->sah_cost = bvh.sah_cost; out->build_ms = (float)bvh.build_ms;
    out->nodes_per_ray = n_rays ? float(double(nv) / n_rays) : 0.0f;
    out->triangles_per_ray = n_rays ? float(double(tt) / n_rays) : 0.0f;
    return MRT_OK;
}
void mrt_noise_fill(uint64_t seed, uint32_t frame, float* tile) { noise_fill(seed, frame, tile); }

}  // extern "C"

// ---- internal: what csrc/noalloc.cu may see of a context (ctx_view.h) ------------------------------------------------
namespace mrt {
bool ctx_view(mrt_ctx* c, CtxView& out) {
    if (!c) return false;
    out.S = c->S; out.stream = c->stream; out.device = c->cfg.device; out.haveScene = c->haveScene;
    out.kernelLaunches = &c->stats.kernel_launches;
    out.image = c->dImage; out.W = c->W; out.H = c->H; out.peerSlot = &c->peer;
    return true;
}
int api_fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    return fail(code, "%s", buf);
}
}  // namespace mrt
