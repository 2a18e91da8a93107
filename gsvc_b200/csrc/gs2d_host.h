// gs2d_host.h -- host-side plumbing shared by the translation units of libgs2d.so.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>

#include "../../include/gs2d.h"

namespace gs2d {

// error text of the calling thread's last failing entry point (capi.cu)
void set_error(const char* fmt, ...);

inline int div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace gs2d

// Every kernel launch is followed by this (the reference checks nothing).
#define GS2D_CHECK_LAUNCH(what)                                                        \
    do {                                                                               \
        cudaError_t e_ = cudaGetLastError();                                           \
        if (e_ != cudaSuccess) {                                                       \
            gs2d::set_error("%s: %s", what, cudaGetErrorString(e_));                   \
            return GS2D_ERR_CUDA;                                                      \
        }                                                                              \
    } while (0)

#define GS2D_CHECK_CUDA(expr)                                                          \
    do {                                                                               \
        cudaError_t e_ = (expr);                                                       \
        if (e_ != cudaSuccess) {                                                       \
            gs2d::set_error("%s: %s", #expr, cudaGetErrorString(e_));                  \
            return GS2D_ERR_CUDA;                                                      \
        }                                                                              \
    } while (0)

#define GS2D_REQUIRE(cond, msg)                                                        \
    do {                                                                               \
        if (!(cond)) {                                                                 \
            gs2d::set_error("%s", msg);                                                \
            return GS2D_ERR_ARG;                                                       \
        }                                                                              \
    } while (0)
