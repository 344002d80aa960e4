// nd_error.h -- thread-local error reporting shared by the translation units.
#pragma once

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "../../include/nanodiloco_b200.h"

namespace nd {

inline char* error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}

inline const char* last_error() { return error_buffer(); }

inline int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

// Launch errors only (cudaPeekAtLastError does not synchronise); execution
// faults surface at the caller's next synchronisation, as with any CUDA work.
inline int check_launch(const char* what) {
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return set_error(ND_ECUDA, "%s: launch failed: %s", what, cudaGetErrorString(e));
    }
    return ND_OK;
}

// NVTX ranges around the phases the library enqueues (K1 / exchange / K2 / AdamW ...), visible in nsys and
// `ncu --nvtx`.  Off unless ND_NVTX=1: a push/pop pair per phase costs nothing next to a multi-GB pass, but a
// library should not emit markers nobody asked for.  nvtx3 is header-only; without a tool attached a call is
// one load and a branch.
inline bool nvtx_enabled() {
    static const bool on = [] { const char* e = getenv("ND_NVTX"); return e && e[0] == '1'; }();
    return on;
}
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char* name) : on(nvtx_enabled()) { if (on) nvtxRangePushA(name); }
    ~NvtxRange() { if (on) nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

#define ND_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return nd::set_error(ND_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_));    \
    } while (0)

}  // namespace nd
