// ctx_view.h — what csrc/noalloc.cu needs to see of a context (mrt_ctx is private to api.cu).
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
};
bool ctx_view(mrt_ctx* c, CtxView& out);      // false if c is NULL (api.cu)
int api_fail(int code, const char* fmt, ...);  // sets mrt_last_error() and returns code (api.cu)

}  // namespace mrt
