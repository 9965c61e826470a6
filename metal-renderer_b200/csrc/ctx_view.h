// ctx_view.h — what csrc/noalloc.cu and csrc/peer.cu need to see of a context (mrt_ctx is private to api.cu).
#pragma once
#include "kernels.h"
#include "mrt.h"

namespace mrt {

struct CtxView {
    DeviceScene S;               // scene + trees as the kernels read them
    cudaStream_t stream;         // the ctx's in-order stream
    int device;
    bool haveScene;
    uint64_t* kernelLaunches;    // mrt_stats.kernel_launches of the ctx
    float4* image;               // accumulation image (null before mrt_resize)
    uint32_t W, H;
    void** peerSlot;             // csrc/peer.cu keeps its per-ctx state here
};
void peer_release(void** slot);               // peer.cu: frees that state (api.cu calls it when the image goes away)
bool ctx_view(mrt_ctx* c, CtxView& out);      // false if c is NULL (api.cu)
int api_fail(int code, const char* fmt, ...);  // sets mrt_last_error() and returns code (api.cu)

}  // namespace mrt
