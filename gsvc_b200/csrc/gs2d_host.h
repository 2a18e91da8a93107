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

// kernel-launch counter and optional per-stage CUDA-event timing (capi.cu); stage ids are
// the GS2D_STAGE_* constants of gs2d.h
void count_launch();
bool prof_enabled();
void prof_begin(int stage, cudaStream_t s);
void prof_end(int stage, cudaStream_t s);

// Brackets the launches of one stage with an event pair on the launching stream when
// profiling is on (gs2d_prof_enable); a single predictable branch when it is off.
// With GS2D_NVTX=1 in the environment every stage is also an NVTX range (named after the stage) on the
// calling thread, so that a timeline profiler shows the library's stages between the caller's own ranges
// (the reference has no tracing hooks at all).
bool nvtx_enabled();
void nvtx_push(int stage);
void nvtx_pop();

struct ProfScope {
    int stage;
    cudaStream_t s;
    bool on, nvtx;
    ProfScope(int stage_, cudaStream_t s_) : stage(stage_), s(s_), on(prof_enabled()), nvtx(nvtx_enabled()) {
        if (nvtx) nvtx_push(stage);
        if (on) prof_begin(stage, s);
    }
    ~ProfScope() {
        if (on) prof_end(stage, s);
        if (nvtx) nvtx_pop();
    }
};

}  // namespace gs2d

// Every kernel launch is followed by this (the reference checks nothing).
#define GS2D_CHECK_LAUNCH(what)                                                        \
    do {                                                                               \
        cudaError_t e_ = cudaGetLastError();                                           \
        if (e_ != cudaSuccess) {                                                       \
            gs2d::set_error("%s: %s", what, cudaGetErrorString(e_));                   \
            return GS2D_ERR_CUDA;                                                      \
        }                                                                              \
        gs2d::count_launch();                                                          \
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
