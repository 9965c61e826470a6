// peer.cu — the framebuffer reduction of sample-sharded rendering done by ONE kernel per GPU over NVLink peer memory.
//
// The reference is single-device (one MTLCommandQueue, Renderer.mm:153); SURVEY.md 8(e) shards samples over GPUs and
// reduces the images.  With NCCL that step is: copy the rank's SUM image, ncclAllReduce, scale by 1 / frames (three
// launches + the collective's own).  Here every rank launches k_peer_reduce on its ctx stream, right behind the kernel
// that finished its image.  Each rank owns an EXCHANGE BLOCK — flags | result image | staging rows — and that block is
// the only memory its peers ever touch (the SUM image itself stays private):
//     push     rank r stores slice g of its image (pixels [g*chunk, (g+1)*chunk)) into staging row r of rank g's block
//              (st.relaxed.sys over NVLink; stores do not wait for a round trip), for every g
//     arrive   the rank's last block to finish pushing stores the step's epoch into its slot of every peer's flags
//              (st.release.sys); every block then polls its OWN flags until all ranks have arrived
//     reduce   rank g sums its N staging rows in rank order 0..N-1 — local loads; every rank and every run gets the same
//              bits — multiplies by `scale`, sets alpha = 1 and stores the pixel into the result image of EVERY rank
//     done     the rank's last block signals "your result slice is written, your staging row is free" to every peer and
//              waits for the same from all of them; only then does the kernel end, so what the stream runs next (a
//              read-back of the result, the next step's pushes) is ordered behind every peer.
// Per rank and step (N ranks, P pixels): P*16 bytes pushed and P*16 bytes of result stored, (N-1)/N of each over NVLink.
//
// How a block reaches the peers: inside one process (the plain-C driver, one thread per GPU) it is allocated with the
// virtual-memory-management driver API and mapped read-write on every GPU of the process — NOT cudaDeviceEnablePeerAccess,
// which maps every allocation of every context into every peer and made a rank's set-up 6-7 s on 8 GPUs (profiles/
// r2_t_*); between processes (torchrun) it is a cudaMalloc allocation opened through CUDA IPC.  mrt_peer_handle
// carries whichever applies.  A wait that sees no progress for about two seconds gives up and sets the rank's error
// word (mrt_peer_status), so a rank that died cannot hang the others.
#include <cstring>
#include <unistd.h>
#include <cuda.h>
#include "ctx_view.h"

namespace mrt {

#define PEER_FLAG_WORDS 64u                 /* arrive[16] | done[16] | error | pushed blocks | reduced blocks | pad: 256 bytes */
#define PEER_ARRIVE 0u
#define PEER_DONE 16u
#define PEER_ERROR 32u
#define PEER_PUSHED 33u
#define PEER_REDUCED 34u
static_assert(MRT_PEER_MAX_RANKS <= 16, "flag block layout");

// exchange block of `pixels` pixels: flags | result[pixels] | staging[world][chunk] with world * chunk <= pixels + 16
static size_t peer_block_bytes(uint32_t pixels) { return PEER_FLAG_WORDS * sizeof(unsigned int) + (2 * (size_t)pixels + MRT_PEER_MAX_RANKS) * sizeof(float4); }

struct PeerArgs {
    unsigned int* block[MRT_PEER_MAX_RANKS];   // every rank's exchange block (own included)
    const float4* image;                       // this rank's SUM image
    uint32_t rank, world, pixels, chunk, epoch;
    float scale;
    long long spinLimit;                       // clock64 ticks a wait may take
};
__device__ __forceinline__ float4* peer_result(unsigned int* block) { return reinterpret_cast<float4*>(block + PEER_FLAG_WORDS); }
__device__ __forceinline__ float4* peer_stage(unsigned int* block, uint32_t pixels) { return peer_result(block) + pixels; }

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
    unsigned int* mine = a.block[a.rank];
    __shared__ unsigned int lastPush, lastReduce;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, step = gridDim.x * blockDim.x;

    // push: slice g of this rank's image -> staging row `rank` of rank g (own slice first, then the peers round-robin)
    for (uint32_t k = 0; k < a.world; ++k) {
        const uint32_t g = (a.rank + k) % a.world;
        const uint32_t lo = min(a.pixels, g * a.chunk), n = min(a.pixels, lo + a.chunk) - lo;
        float4* dst = peer_stage(a.block[g], a.pixels) + (size_t)a.rank * a.chunk;
        for (uint32_t i = tid; i < n; i += step) st_relaxed_sys(dst + i, a.image[lo + i]);
    }
    __threadfence_system();                                 // this thread's pushes are visible system-wide ...
    __syncthreads();                                        // ... for every thread of the block
    if (threadIdx.x == 0) lastPush = (atomicAdd(mine + PEER_PUSHED, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (lastPush && threadIdx.x < a.world) {                // the whole rank has pushed: tell everybody
        __threadfence_system();
        st_release_sys(a.block[threadIdx.x] + PEER_ARRIVE + a.rank, a.epoch);
    }
    if (threadIdx.x < a.world && !wait_for(mine + PEER_ARRIVE + threadIdx.x, a.epoch, a.spinLimit)) atomicExch(mine + PEER_ERROR, 1u + threadIdx.x);
    __syncthreads();

    // reduce this rank's slice out of its own staging rows, publish it to every rank's result image
    const uint32_t lo = min(a.pixels, a.rank * a.chunk), n = min(a.pixels, lo + a.chunk) - lo;
    const float4* rows = peer_stage(mine, a.pixels);
    for (uint32_t i = tid; i < n; i += step) {
        float4 s = ld_relaxed_sys(rows + i);                // written by a peer: not through this SM's L1
        for (uint32_t r = 1; r < a.world; ++r) {
            const float4 v = ld_relaxed_sys(rows + (size_t)r * a.chunk + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w = 1.0f;          // alpha = 1, like k_scale_image
        for (uint32_t k = 0; k < a.world; ++k) st_relaxed_sys(peer_result(a.block[(a.rank + k) % a.world]) + lo + i, s);
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) lastReduce = (atomicAdd(mine + PEER_REDUCED, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (!lastReduce) return;
    if (threadIdx.x == 0) { mine[PEER_PUSHED] = 0u; mine[PEER_REDUCED] = 0u; }   // ready for the next step (nobody else touches them)
    __threadfence_system();
    if (threadIdx.x < a.world) {
        st_release_sys(a.block[threadIdx.x] + PEER_DONE + a.rank, a.epoch);
        if (!wait_for(mine + PEER_DONE + threadIdx.x, a.epoch, a.spinLimit)) atomicExch(mine + PEER_ERROR, 0x100u + threadIdx.x);
    }
}

// ---- host side -------------------------------------------------------------------------------------------------
// Driver entry points of the virtual-memory-management API, looked up through the runtime (the library links no libcuda).
struct Vmm {
    CUresult (*getGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*setAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*addressFree)(CUdeviceptr, size_t) = nullptr;
    bool ok = false;
};
static const Vmm& vmm() {
    static const Vmm v = [] {
        Vmm t;
        auto get = [](const char* name, void** fn) {
            cudaDriverEntryPointQueryResult q;
            return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
        };
        t.ok = get("cuMemGetAllocationGranularity", (void**)&t.getGranularity) && get("cuMemCreate", (void**)&t.create) &&
               get("cuMemAddressReserve", (void**)&t.reserve) && get("cuMemMap", (void**)&t.map) && get("cuMemSetAccess", (void**)&t.setAccess) &&
               get("cuMemUnmap", (void**)&t.unmap) && get("cuMemRelease", (void**)&t.release) && get("cuMemAddressFree", (void**)&t.addressFree);
        cudaGetLastError();
        return t;
    }();
    return v;
}

struct PeerState {
    void* block = nullptr;                  // this ctx's exchange block
    size_t blockBytes = 0, mappedBytes = 0; // mappedBytes != 0: a VMM mapping (shared inside the process), else cudaMalloc (IPC)
    CUmemGenericAllocationHandle vmmHandle = 0;
    uint32_t pixels = 0, rank = 0, world = 0, epoch = 0;
    bool connected = false;
    PeerArgs args{};
    void* opened[MRT_PEER_MAX_RANKS] = {};  // IPC mappings to close
    int nOpened = 0;
};

static PeerState* state_of(const CtxView& v) { return static_cast<PeerState*>(*v.peerSlot); }

// Block shared inside the process: physical memory on `device`, mapped read-write on every GPU that can reach it.
static bool vmm_alloc(PeerState* st, int device, size_t bytes) {
    const Vmm& m = vmm();
    if (!m.ok) return false;
    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id = device;
    size_t gran = 0;
    if (m.getGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0) return false;
    const size_t size = (bytes + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle h = 0;
    if (m.create(&h, size, &prop, 0) != CUDA_SUCCESS) return false;
    CUdeviceptr ptr = 0;
    if (m.reserve(&ptr, size, 0, 0, 0) != CUDA_SUCCESS) { m.release(h); return false; }
    if (m.map(ptr, size, 0, h, 0) != CUDA_SUCCESS) { m.addressFree(ptr, size); m.release(h); return false; }
    int count = 0;
    cudaGetDeviceCount(&count);
    CUmemAccessDesc acc[64];
    size_t n = 0;
    for (int d = 0; d < count && n < 64; ++d) {
        int can = d == device;
        if (!can) cudaDeviceCanAccessPeer(&can, d, device);
        if (!can) continue;
        acc[n].location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc[n].location.id = d; acc[n].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; ++n;
    }
    if (m.setAccess(ptr, size, acc, n) != CUDA_SUCCESS) { m.unmap(ptr, size); m.addressFree(ptr, size); m.release(h); return false; }
    st->block = (void*)(uintptr_t)ptr; st->mappedBytes = size; st->vmmHandle = h;
    return true;
}

static void peer_close(PeerState* st) {
    for (int i = 0; i < st->nOpened; ++i) cudaIpcCloseMemHandle(st->opened[i]);
    st->nOpened = 0; st->connected = false;
}

void peer_release(void** slot) {                            // called by api.cu when the image is freed (resize, destroy)
    PeerState* st = static_cast<PeerState*>(*slot);
    if (!st) return;
    peer_close(st);
    if (st->mappedBytes) {
        const Vmm& m = vmm();
        const CUdeviceptr p = (CUdeviceptr)(uintptr_t)st->block;
        m.unmap(p, st->mappedBytes); m.addressFree(p, st->mappedBytes); m.release(st->vmmHandle);
    } else {
        cudaFree(st->block);
    }
    delete st;
    *slot = nullptr;
}

}  // namespace mrt

using namespace mrt;

extern "C" {

int mrt_peer_export(mrt_ctx* ctx, uint32_t sharing, mrt_peer_handle* out) {
    CtxView v;
    if (!ctx_view(ctx, v) || !out) return api_fail(MRT_E_INVALID, "mrt_peer_export: NULL argument");
    if (sharing > MRT_PEER_SHARE_PROCESS) return api_fail(MRT_E_INVALID, "mrt_peer_export: sharing must be MRT_PEER_SHARE_IPC or MRT_PEER_SHARE_PROCESS");
    if (!v.image) return api_fail(MRT_E_STATE, "no size: call mrt_resize first");
    cudaError_t e = cudaSetDevice(v.device);
    if (e == cudaSuccess) e = cudaStreamSynchronize(v.stream);
    if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "mrt_peer_export: %s", cudaGetErrorString(e));
    PeerState* st = state_of(v);
    // An existing block of the same kind is handed out again: another process may still have it open through IPC, and
    // freeing exported memory under an importer is undefined.  (A resize frees it: disconnect the peers first.)
    if (st && (st->pixels != v.W * v.H || (st->mappedBytes != 0) != (sharing == MRT_PEER_SHARE_PROCESS && vmm().ok))) { peer_release(v.peerSlot); st = nullptr; }
    if (!st) {
        st = new PeerState();
        *v.peerSlot = st;
        st->pixels = v.W * v.H;
        st->blockBytes = peer_block_bytes(st->pixels);
        if (sharing == MRT_PEER_SHARE_PROCESS) vmm_alloc(st, v.device, st->blockBytes);      // an old driver falls through to a plain
        if (!st->block) e = cudaMalloc(&st->block, st->blockBytes);                          // allocation: peers then need peer access
        if (e == cudaSuccess) e = cudaMemset(st->block, 0, st->blockBytes);
        if (e != cudaSuccess) { peer_release(v.peerSlot); return api_fail(MRT_E_CUDA, "mrt_peer_export: %s", cudaGetErrorString(e)); }
    }
    std::memset(out, 0, sizeof *out);
    out->block_ptr = (uint64_t)(uintptr_t)st->block;
    out->pid = (int64_t)getpid(); out->device = v.device; out->width = v.W; out->height = v.H;
    out->mapped_everywhere = st->mappedBytes ? 1u : 0u;
    static_assert(sizeof(cudaIpcMemHandle_t) == sizeof out->block_ipc, "IPC handle size");
    cudaIpcMemHandle_t h;
    if (!st->mappedBytes && cudaIpcGetMemHandle(&h, st->block) == cudaSuccess) { std::memcpy(out->block_ipc, &h, sizeof h); out->ipc_valid = 1; }
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
        void* block = nullptr;
        if (h.pid == (int64_t)getpid()) {                   // same process: the pointer itself
            block = (void*)(uintptr_t)h.block_ptr;
            if (h.device != v.device && !h.mapped_everywhere) {     // a plain allocation on another GPU: whole-context peer access
                int can = 0;
                cudaDeviceCanAccessPeer(&can, v.device, h.device);
                if (!can) return api_fail(MRT_E_STATE, "mrt_peer_connect: device %d cannot map the memory of device %d", v.device, h.device);
                e = cudaDeviceEnablePeerAccess(h.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
                if (e != cudaSuccess) return api_fail(MRT_E_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", h.device, cudaGetErrorString(e));
            }
        } else {                                            // another process: CUDA IPC
            if (!h.ipc_valid) return api_fail(MRT_E_STATE, "mrt_peer_connect: rank %u (another process) exported no IPC handle — export with MRT_PEER_SHARE_IPC", r);
            cudaIpcMemHandle_t ih;
            std::memcpy(&ih, h.block_ipc, sizeof ih);
            e = cudaIpcOpenMemHandle(&block, ih, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { peer_close(st); return api_fail(MRT_E_CUDA, "cudaIpcOpenMemHandle (rank %u): %s", r, cudaGetErrorString(e)); }
            st->opened[st->nOpened++] = block;
        }
        a.block[r] = static_cast<unsigned int*>(block);
    }
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, v.device);
    a.image = v.image; a.rank = rank; a.world = world; a.pixels = st->pixels; a.chunk = (st->pixels + world - 1u) / world;
    a.spinLimit = 2ll * 1000ll * (long long)(khz > 0 ? khz : 2000000);   // ~2 s
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
    // few enough blocks that the kernels of every rank are resident at once even when several ranks share one GPU (tests):
    // every block waits for every rank, so a block that cannot be scheduled would stall them all
    uint32_t blocks = (st->pixels + 1023u) / 1024u;
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
                        (err & 0x100u) ? "finish" : "arrive", st->epoch);
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
    if (state_of(v)) peer_close(state_of(v));               // this rank's mappings of its peers' blocks; its own block stays
    return MRT_OK;
}

}  // extern "C"
