// noalloc.cu — mrt_intersect_host_noalloc: the stand-alone nearest-hit pass of mrt_intersect_host (csrc/api.cu) for callers
// that must not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
//
// mrt_intersect_host cudaMallocs two scratch buffers per call (the staged rays and the hit records) and frees them.
// Here that scratch — plus the work counter of the persistent traversal — lives in ONE file-scope
//     static __device__ __align__(256) unsigned char g_noalloc_scratch[MRT_NOALLOC_SCRATCH_BYTES]      (6 MiB + 256 B)
// which every GPU that loads the library holds until the process exits.  Layout:
//     [0, 256)                      work counter of the persistent kernel (word 0), cleared to zero by every call
//     [256, 256 + 65,536 * 80)      the caller's ray records, copied verbatim (count * stride bytes)
//     [.., + 65,536 * 16)           the hit records the kernel writes (every ray writes its own, so nothing to clear)
// The name of the array is a HOST-SIDE STAND-IN in host code, not a device address: its address is resolved once per
// device, with that device current, by cudaGetSymbolAddress (checked), and that pointer is what the copies, the memset and
// the kernel receive.  The call uses cudaMemsetAsync / cudaMemcpyAsync / one kernel launch / cudaStreamSynchronize on the
// ctx stream and nothing that allocates.
//
// Every call on a device shares the array, and it keeps the last call's contents: calls that overlap on the GPU (two
// contexts, two threads, two processes' worth of streams on one device) RACE — serialise them.
#include <mutex>
#include "ctx_view.h"
#include "device_trace8.cuh"

namespace mrt {

#define NOALLOC_HEAD_BYTES 256u
#define NOALLOC_RAY_BYTES ((size_t)MRT_NOALLOC_MAX_RAYS * MRT_NOALLOC_MAX_STRIDE)
#define NOALLOC_HIT_BYTES ((size_t)MRT_NOALLOC_MAX_RAYS * sizeof(mrt_intersection))
static_assert(MRT_NOALLOC_SCRATCH_BYTES == NOALLOC_HEAD_BYTES + NOALLOC_RAY_BYTES + NOALLOC_HIT_BYTES, "scratch layout");
static_assert(NOALLOC_RAY_BYTES % 256 == 0, "hit records start 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[MRT_NOALLOC_SCRATCH_BYTES];

// Device address of the array on each device, resolved on the first call there (set-up, once per device).
#define NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mutex;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

static cudaError_t noalloc_scratch(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        unsigned char* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);     // `device` is current (caller set it)
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

// IO policy of the persistent traversal over the staged rays: same record convention as the MPS encode
// (Renderer.mm:464-469, 519-523): origin[3], minDistance, direction[3], maxDistance; maxDistance < 0 = skipped = miss.
struct ScratchIO {
    const unsigned char* rays; uint32_t stride; bool vec; mrt_intersection* out;
    __device__ __forceinline__ bool load(uint32_t i, TraceJob& job) {
        const unsigned char* p = rays + (size_t)i * stride;
        float4 a, b;
        if (vec) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 16); }
        else {
            const float* f = reinterpret_cast<const float*>(p);
            a = make_float4(f[0], f[1], f[2], f[3]); b = make_float4(f[4], f[5], f[6], f[7]);
        }
        if (!(b.w >= 0.0f)) {
            *reinterpret_cast<float4*>(out + i) = make_float4(-1.0f, __uint_as_float(MRT_NO_HIT), 0.0f, 0.0f);
            return false;
        }
        job.o = mk3(a.x, a.y, a.z); job.d = mk3(b.x, b.y, b.z); job.tmin = a.w; job.tmax = b.w;
        job.target = MRT_NO_HIT; job.tag = i;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, uint32_t, const HitRec& h) {
        *reinterpret_cast<float4*>(out + i) = make_float4(h.t, __uint_as_float(h.tri), h.u, h.v);
    }
};

// The production traversal core (trace_persistent8: 8-wide quantised tree, vote scheduling, dynamic refill) over the
// scratch array: work counter, rays and hit records all live behind `scratch`.
__global__ void __launch_bounds__(128, 8) k_intersect_noalloc(DeviceScene S, unsigned char* scratch, uint32_t stride, uint32_t count,
                                                              bool vec, uint32_t refillBelow, uint32_t leafMin, uint32_t magic) {
    __shared__ __align__(16) uint8_t orderLut[MRT_ORDER_LUT_BYTES];
    bvh8_fill_order_lut(orderLut);
    ScratchIO io;
    io.rays = scratch + NOALLOC_HEAD_BYTES; io.stride = stride; io.vec = vec;
    io.out = reinterpret_cast<mrt_intersection*>(scratch + NOALLOC_HEAD_BYTES + NOALLOC_RAY_BYTES);
    trace_persistent8(S.nodes8, S.tris8, count, reinterpret_cast<unsigned int*>(scratch), refillBelow, leafMin, magic, orderLut, io);
}

}  // namespace mrt

using namespace mrt;

extern "C" int mrt_intersect_host_noalloc(mrt_ctx* ctx, const void* rays, size_t stride_bytes, uint32_t count, mrt_intersection* out) {
    CtxView v;
    if (!ctx_view(ctx, v) || !rays || !out) return api_fail(MRT_E_INVALID, "mrt_intersect_host_noalloc: NULL argument");
    if (!v.haveScene) return api_fail(MRT_E_STATE, "no scene: call mrt_set_scene first");
    if (stride_bytes < 32 || stride_bytes % 4 != 0) return api_fail(MRT_E_INVALID, "ray stride must be >= 32 and a multiple of 4");
    if (stride_bytes > MRT_NOALLOC_MAX_STRIDE)
        return api_fail(MRT_E_INVALID, "mrt_intersect_host_noalloc: stride %zu exceeds the static scratch's maximum of %u bytes per ray "
                        "(use mrt_intersect_host)", stride_bytes, (unsigned)MRT_NOALLOC_MAX_STRIDE);
    if (count > MRT_NOALLOC_MAX_RAYS)
        return api_fail(MRT_E_INVALID, "mrt_intersect_host_noalloc: %u rays exceed the static scratch's maximum of %u rays per call "
                        "(split the batch or use mrt_intersect_host)", count, (unsigned)MRT_NOALLOC_MAX_RAYS);
    if (v.device < 0 || v.device >= NOALLOC_MAX_DEVICES) return api_fail(MRT_E_INVALID, "device %d out of range for the no-alloc path", v.device);
    if (count == 0) return MRT_OK;
    cudaError_t e = cudaSetDevice(v.device);
    unsigned char* scratch = nullptr;
    if (e == cudaSuccess) e = noalloc_scratch(v.device, &scratch);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_intersect_host_noalloc: cudaGetSymbolAddress of the scratch array failed: %s", cudaGetErrorString(e));
    const size_t rayBytes = (size_t)count * stride_bytes;
    unsigned char* dRays = scratch + NOALLOC_HEAD_BYTES;
    unsigned char* dHits = scratch + NOALLOC_HEAD_BYTES + NOALLOC_RAY_BYTES;
    e = cudaMemsetAsync(scratch, 0, NOALLOC_HEAD_BYTES, v.stream);                      // the work counter must start at zero
    if (e == cudaSuccess) e = cudaMemcpyAsync(dRays, rays, rayBytes, cudaMemcpyDefault, v.stream);   // host or device rays
    if (e == cudaSuccess) {
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, v.device);
        const uint32_t cap = (uint32_t)(sms > 0 ? sms : 148) * 8u, need = ((count + 31u) / 32u + 3u) / 4u;
        const bool vec = stride_bytes % 16 == 0;                                         // dRays is 256-byte aligned
        // vote tuning of the production trace kernel (csrc/wavefront.cu): refill below 24 lanes, 10 lanes / ratio 7 of 16
        k_intersect_noalloc<<<need < cap ? need : cap, 128, 0, v.stream>>>(v.S, scratch, (uint32_t)stride_bytes, count, vec, 24u,
                                                                           10u | (7u << 8) | (16u << 18), 0x47000000u);
        *v.kernelLaunches += 1;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dHits, (size_t)count * sizeof(mrt_intersection), cudaMemcpyDefault, v.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(v.stream);                          // rays / out are borrowed for the call only
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_intersect_host_noalloc: %s", cudaGetErrorString(e));
    return MRT_OK;
}
