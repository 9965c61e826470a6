// peer.cu — the framebuffer reduction of sample-sharded rendering done by ONE kernel per GPU over NVLink peer memory.
//
// The reference is single-device (one MTLCommandQueue, Renderer.mm:153); SURVEY.md 8(e) shards samples over GPUs and
// reduces the images.  With NCCL that step is: copy the rank's SUM image, ncclAllReduce, scale by 1 / frames (three
// launches + the collective's own).  Here every rank launches k_peer_reduce on its ctx stream, right behind the kernel
// that finished its image:
//     arrive   block 0 stores the step's epoch into its slot of every peer's flag block (st.release.sys)
//     wait     every block polls its OWN flag block until all peers have arrived (their images are complete)
//     reduce   rank g sums pixels [g*chunk, (g+1)*chunk) of all N images straight out of the peers' memory
//              (ld.relaxed.sys, rank order 0..N-1 — every rank and every run gets the same bits), multiplies by
//              `scale` and stores the pixel into the result buffer of EVERY rank (reduce-scatter + all-gather in one pass)
//     done     the last block of the rank signals "I have read your image and written your result" to every peer and
//              waits for the same from all of them; only then does the kernel end, so whatever the stream runs next (the
//              next batch's accumulate into the SUM image, a read-back of the result) is ordered behind every peer's use.
// Per rank and step (N ranks, P pixels): P*16 bytes read (P/N of them local), P*16 written, two flag rounds.
//
// Buffers reach the peers as raw pointers inside one process (cudaDeviceEnablePeerAccess; the plain-C driver runs one
// thread per GPU) or as CUDA IPC handles between processes (torchrun); mrt_peer_handle carries both.  A wait that sees
// no progress for about two seconds gives up and sets the rank's error word (mrt_peer_status), so a rank that died cannot
// hang the others.
#include <cstring>
#include <unistd.h>
#include "ctx_view.h"

namespace mrt {

#define PEER_FLAG_WORDS 64u                 /* arrive[16] | done[16] | error | blocks finished | pad: 256 bytes */
#define PEER_ARRIVE 0u
#define PEER_DONE 16u
#define PEER_ERROR 32u
#define PEER_BLOCKS 33u
static_assert(MRT_PEER_MAX_RANKS <= 16, "flag block layout");

struct PeerArgs {
    const float4* image[MRT_PEER_MAX_RANKS];
    float4* result[MRT_PEER_MAX_RANKS];
    unsigned int* flags[MRT_PEER_MAX_RANKS];
    uint32_t rank, world, pixels, epoch;
    float scale;
    long long spinLimit;                    // clock64 ticks a wait may take
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(float4* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// true when flag word `p` has reached `epoch` (epochs only grow; compared as a signed distance so that wrap-around is harmless)
__device__ __forceinline__ bool wait_for(const unsigned int* p, unsigned int epoch, long long limit) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(p) - epoch) < 0) {
        if (clock64() - t0 > limit) return false;
        __nanosleep(64);
    }
    return true;
}

__global__ void __launch_bounds__(256) k_peer_reduce(PeerArgs a) {
    unsigned int* mine = a.flags[a.rank];
    __shared__ unsigned int last;
    if (blockIdx.x == 0 && threadIdx.x < a.world) st_release_sys(a.flags[threadIdx.x] + PEER_ARRIVE + a.rank, a.epoch);
    if (threadIdx.x < a.world && !wait_for(mine + PEER_ARRIVE + threadIdx.x, a.epoch, a.spinLimit)) atomicExch(mine + PEER_ERROR, 1u + threadIdx.x);
    __syncthreads();

    const uint32_t chunk = (a.pixels + a.world - 1u) / a.world;
    const uint32_t lo = min(a.pixels, a.rank * chunk), hi = min(a.pixels, lo + chunk);
    for (uint32_t i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) {
        float4 s = ld_relaxed_sys(a.image[0] + i);
        for (uint32_t r = 1; r < a.world; ++r) {
            const float4 v = ld_relaxed_sys(a.image[r] + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w = 1.0f;          // alpha = 1, like k_scale_image
        for (uint32_t r = 0; r < a.world; ++r) st_relaxed_sys(a.result[r] + i, s);
    }

    __threadfence_system();                                 // this thread's result stores are visible system-wide ...
    __syncthreads();                                        // ... for every thread of the block
    if (threadIdx.x == 0) last = (atomicAdd(mine + PEER_BLOCKS, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (!last) return;
    if (threadIdx.x == 0) mine[PEER_BLOCKS] = 0u;           // ready for the next step (nobody else touches it)
    __threadfence_system();
    if (threadIdx.x < a.world) {
        st_release_sys(a.flags[threadIdx.x] + PEER_DONE + a.rank, a.epoch);
        if (!wait_for(mine + PEER_DONE + threadIdx.x, a.epoch, a.spinLimit)) atomicExch(mine + PEER_ERROR, 0x100u + threadIdx.x);
    }
}

// ---- host side -------------------------------------------------------------------------------------------------
struct PeerState {
    void* block = nullptr;                  // this ctx's flag block + result buffer (one cudaMalloc)
    size_t blockBytes = 0;
    uint32_t pixels = 0, rank = 0, world = 0, epoch = 0;
    bool connected = false;
    PeerArgs args{};
    void* opened[2 * MRT_PEER_MAX_RANKS] = {};              // IPC mappings to close
    int nOpened = 0;
};

static PeerState* state_of(const CtxView& v) { return static_cast<PeerState*>(*v.peerSlot); }

static void peer_close(PeerState* st) {
    for (int i = 0; i < st->nOpened; ++i) cudaIpcCloseMemHandle(st->opened[i]);
    st->nOpened = 0; st->connected = false;
}

void peer_release(void** slot) {                            // called by api.cu when the image is freed (resize, destroy)
    PeerState* st = static_cast<PeerState*>(*slot);
    if (!st) return;
    peer_close(st);
    cudaFree(st->block);
    delete st;
    *slot = nullptr;
}

}  // namespace mrt

using namespace mrt;

extern "C" {

int mrt_peer_export(mrt_ctx* ctx, mrt_peer_handle* out) {
    CtxView v;
    if (!ctx_view(ctx, v) || !out) return api_fail(MRT_E_INVALID, "mrt_peer_export: NULL argument");
    if (!v.image) return api_fail(MRT_E_STATE, "no size: call mrt_resize first");
    cudaError_t e = cudaSetDevice(v.device);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    PeerState* st = state_of(v);
    const uint32_t pixels = v.W * v.H;
    if (!st) { st = new PeerState(); *v.peerSlot = st; }
    if (!st->block) {
        st->pixels = pixels;
        st->blockBytes = PEER_FLAG_WORDS * sizeof(unsigned int) + (size_t)pixels * sizeof(float4);
        e = cudaMalloc(&st->block, st->blockBytes);
        if (e == cudaSuccess) e = cudaMemset(st->block, 0, st->blockBytes);
        if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_export: %s", cudaGetErrorString(e));
    }
    std::memset(out, 0, sizeof *out);
    out->image_ptr = (uint64_t)(uintptr_t)v.image; out->block_ptr = (uint64_t)(uintptr_t)st->block;
    out->pid = (int64_t)getpid(); out->device = v.device; out->width = v.W; out->height = v.H;
    static_assert(sizeof(cudaIpcMemHandle_t) == sizeof out->image_ipc, "IPC handle size");
    cudaIpcMemHandle_t h;
    out->ipc_valid = 0;
    if (cudaIpcGetMemHandle(&h, v.image) == cudaSuccess) {
        std::memcpy(out->image_ipc, &h, sizeof h);
        if (cudaIpcGetMemHandle(&h, st->block) == cudaSuccess) { std::memcpy(out->block_ipc, &h, sizeof h); out->ipc_valid = 1; }
    }
    cudaGetLastError();                                     // a platform without IPC still works inside one process
    return MRT_OK;
}

int mrt_peer_connect(mrt_ctx* ctx, uint32_t rank, uint32_t world, const mrt_peer_handle* handles) {
    CtxView v;
    if (!ctx_view(ctx, v) || !handles) return api_fail(MRT_E_INVALID, "mrt_peer_connect: NULL argument");
    if (world < 1 || world > MRT_PEER_MAX_RANKS || rank >= world) return api_fail(MRT_E_INVALID, "mrt_peer_connect: need rank < world <= %d", MRT_PEER_MAX_RANKS);
    PeerState* st = state_of(v);
    if (!st || !st->block) return api_fail(MRT_E_STATE, "mrt_peer_connect: call mrt_peer_export first");
    if (handles[rank].block_ptr != (uint64_t)(uintptr_t)st->block || handles[rank].pid != (int64_t)getpid())
        return api_fail(MRT_E_INVALID, "mrt_peer_connect: handles[rank] is not this context's handle");
    cudaError_t e = cudaSetDevice(v.device);
    if (e == cudaSuccess) e = cudaStreamSynchronize(v.stream);      // an earlier group's last reduction has left the flags alone
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_connect: %s", cudaGetErrorString(e));
    peer_close(st);
    PeerArgs& a = st->args;
    a = PeerArgs{};
    for (uint32_t r = 0; r < world; ++r) {
        const mrt_peer_handle& h = handles[r];
        if (h.width != v.W || h.height != v.H) return api_fail(MRT_E_INVALID, "mrt_peer_connect: rank %u renders %ux%u, this one %ux%u", r, h.width, h.height, v.W, v.H);
        void* image = nullptr; void* block = nullptr;
        if (h.pid == (int64_t)getpid()) {                   // same process: raw pointers, peer access between the two devices
            image = (void*)(uintptr_t)h.image_ptr; block = (void*)(uintptr_t)h.block_ptr;
            if (h.device != v.device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, v.device, h.device);
                if (!can) return api_fail(MRT_E_STATE, "mrt_peer_connect: device %d cannot map the memory of device %d", v.device, h.device);
                e = cudaDeviceEnablePeerAccess(h.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
                if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", h.device, cudaGetErrorString(e));
            }
        } else {                                            // another process: CUDA IPC
            if (!h.ipc_valid) return api_fail(MRT_E_STATE, "mrt_peer_connect: rank %u exported no IPC handles", r);
            cudaIpcMemHandle_t ih;
            std::memcpy(&ih, h.image_ipc, sizeof ih);
            e = cudaIpcOpenMemHandle(&image, ih, cudaIpcMemLazyEnablePeerAccess);
            if (e == cudaSuccess) { st->opened[st->nOpened++] = image; std::memcpy(&ih, h.block_ipc, sizeof ih); e = cudaIpcOpenMemHandle(&block, ih, cudaIpcMemLazyEnablePeerAccess); }
            if (e == cudaSuccess) st->opened[st->nOpened++] = block;
            if (e != cudaSuccess) { peer_close(st); return api_fail(MRT_E_CUDA, "cudaIpcOpenMemHandle (rank %u): %s", r, cudaGetErrorString(e)); }
        }
        a.image[r] = static_cast<const float4*>(image);
        a.flags[r] = static_cast<unsigned int*>(block);
        a.result[r] = reinterpret_cast<float4*>(static_cast<unsigned int*>(block) + PEER_FLAG_WORDS);
    }
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, v.device);
    a.rank = rank; a.world = world; a.pixels = st->pixels; a.spinLimit = 2ll * 1000ll * (long long)(khz > 0 ? khz : 2000000);   // ~2 s
    st->rank = rank; st->world = world; st->epoch = 0; st->connected = true;
    e = cudaMemset(st->block, 0, PEER_FLAG_WORDS * sizeof(unsigned int));
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_connect: %s", cudaGetErrorString(e));
    return MRT_OK;
}

int mrt_peer_reduce(mrt_ctx* ctx, float scale) {
    CtxView v;
    if (!ctx_view(ctx, v)) return api_fail(MRT_E_INVALID, "ctx is NULL");
    PeerState* st = state_of(v);
    if (!st || !st->connected) return api_fail(MRT_E_STATE, "mrt_peer_reduce: call mrt_peer_connect first (and again after mrt_resize)");
    cudaError_t e = cudaSetDevice(v.device);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    st->args.epoch = ++st->epoch; st->args.scale = scale;
    // few enough blocks that the kernels of every rank are resident at once even when several ranks share one GPU (tests)
    const uint32_t chunk = (st->pixels + st->world - 1u) / st->world;
    uint32_t blocks = (chunk + 1023u) / 1024u;
    const uint32_t cap = st->world <= 8u ? 96u : 48u;       // world * cap blocks of 256 threads fit one GPU (768 of >= 888)
    if (blocks > cap) blocks = cap;
    if (blocks < 1u) blocks = 1u;
    k_peer_reduce<<<blocks, 256, 0, v.stream>>>(st->args);
    *v.kernelLaunches += 1;
    e = cudaGetLastError();
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "k_peer_reduce: %s", cudaGetErrorString(e));
    return MRT_OK;
}

int mrt_peer_result(mrt_ctx* ctx, void** device_ptr, size_t* bytes) {
    CtxView v;
    if (!ctx_view(ctx, v)) return api_fail(MRT_E_INVALID, "ctx is NULL");
    PeerState* st = state_of(v);
    if (!st || !st->block) return api_fail(MRT_E_STATE, "mrt_peer_result: call mrt_peer_export first");
    if (device_ptr) *device_ptr = static_cast<unsigned int*>(st->block) + PEER_FLAG_WORDS;
    if (bytes) *bytes = (size_t)st->pixels * sizeof(float4);
    return MRT_OK;
}

int mrt_peer_status(mrt_ctx* ctx) {
    CtxView v;
    if (!ctx_view(ctx, v)) return api_fail(MRT_E_INVALID, "ctx is NULL");
    PeerState* st = state_of(v);
    if (!st || !st->block) return MRT_OK;
    cudaError_t e = cudaSetDevice(v.device);
    if (e == cudaSuccess) e = cudaStreamSynchronize(v.stream);
    unsigned int err = 0;
    if (e == cudaSuccess) e = cudaMemcpy(&err, static_cast<unsigned int*>(st->block) + PEER_ERROR, sizeof err, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_status: %s", cudaGetErrorString(e));
    if (err) {
        cudaMemset(static_cast<unsigned int*>(st->block) + PEER_ERROR, 0, sizeof err);
        return api_fail(MRT_E_STATE, "peer reduction: rank %u gave up waiting for rank %u to %s (step %u)", st->rank, (err & 0xffu) - ((err & 0x100u) ? 0u : 1u),
                        (err & 0x100u) ? "finish reading" : "arrive", st->epoch);
    }
    return MRT_OK;
}

int mrt_peer_read_result(mrt_ctx* ctx, float* rgba) {
    CtxView v;
    if (!ctx_view(ctx, v) || !rgba) return api_fail(MRT_E_INVALID, "mrt_peer_read_result: NULL argument");
    int rc = mrt_peer_status(ctx);                           // synchronises; reports a wait that gave up
    if (rc != MRT_OK) return rc;
    PeerState* st = state_of(v);
    if (!st || !st->block) return api_fail(MRT_E_STATE, "mrt_peer_read_result: call mrt_peer_export first");
    cudaError_t e = cudaMemcpy(rgba, static_cast<unsigned int*>(st->block) + PEER_FLAG_WORDS, (size_t)st->pixels * sizeof(float4), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_read_result: %s", cudaGetErrorString(e));
    return MRT_OK;
}

int mrt_peer_disconnect(mrt_ctx* ctx) {
    CtxView v;
    if (!ctx_view(ctx, v)) return api_fail(MRT_E_INVALID, "ctx is NULL");
    if (cudaSetDevice(v.device) == cudaSuccess) cudaStreamSynchronize(v.stream);
    peer_release(v.peerSlot);
    return MRT_OK;
}

}  // extern "C"
