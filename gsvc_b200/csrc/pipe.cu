// pipe.cu -- frame pipe: raw frames from pinned host memory into a ring of device slots on a
// copy stream of its own, so that the copy of frame i+1 runs under the fit step of frame i.
// Pure host-side stream/event plumbing (six driver calls per frame); it exists natively because
// the same plumbing through a Python tensor library costs more host time per frame than the
// whole fit step takes on the device.
#include <cstdint>
#include <new>

#include "gs2d_host.h"

struct gs2d_frame_pipe {
    char* slots;
    size_t frame_bytes;
    int depth;
    int next;
    cudaStream_t copy_stream;
    cudaEvent_t copied[GS2D_PIPE_MAX_DEPTH];
    cudaEvent_t freed[GS2D_PIPE_MAX_DEPTH];
    bool has_consumer[GS2D_PIPE_MAX_DEPTH];
};

using namespace gs2d;

extern "C" {

int gs2d_frame_pipe_create(void* device_slots, size_t frame_bytes, int depth, gs2d_frame_pipe** out) {
    GS2D_REQUIRE(out != nullptr, "frame_pipe_create: null out");
    *out = nullptr;
    GS2D_REQUIRE(device_slots != nullptr && frame_bytes > 0, "frame_pipe_create: no slots");
    GS2D_REQUIRE(depth >= 1 && depth <= GS2D_PIPE_MAX_DEPTH, "frame_pipe_create: depth out of range");
    gs2d_frame_pipe* p = new (std::nothrow) gs2d_frame_pipe();
    GS2D_REQUIRE(p != nullptr, "frame_pipe_create: out of host memory");
    p->slots = (char*)device_slots;
    p->frame_bytes = frame_bytes;
    p->depth = depth;
    p->next = 0;
    cudaError_t e = cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < depth && e == cudaSuccess; ++i) {
        p->has_consumer[i] = false;
        e = cudaEventCreateWithFlags(&p->copied[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->freed[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        set_error("frame_pipe_create: %s", cudaGetErrorString(e));
        delete p;   // the context is unusable anyway
        return GS2D_ERR_CUDA;
    }
    *out = p;
    return GS2D_OK;
}

int gs2d_frame_pipe_destroy(gs2d_frame_pipe* p) {
    if (!p) return GS2D_OK;
    cudaStreamSynchronize(p->copy_stream);
    for (int i = 0; i < p->depth; ++i) {
        cudaEventDestroy(p->copied[i]);
        cudaEventDestroy(p->freed[i]);
    }
    cudaStreamDestroy(p->copy_stream);
    delete p;
    return GS2D_OK;
}

int gs2d_frame_pipe_push(gs2d_frame_pipe* p, const void* host_frame, int* slot_out) {
    GS2D_REQUIRE(p && host_frame && slot_out, "frame_pipe_push: null argument");
    const int b = p->next;
    p->next = (b + 1) % p->depth;
    if (p->has_consumer[b]) GS2D_CHECK_CUDA(cudaStreamWaitEvent(p->copy_stream, p->freed[b], 0));
    GS2D_CHECK_CUDA(cudaMemcpyAsync(p->slots + (size_t)b * p->frame_bytes, host_frame, p->frame_bytes,
                                    cudaMemcpyHostToDevice, p->copy_stream));
    GS2D_CHECK_CUDA(cudaEventRecord(p->copied[b], p->copy_stream));
    *slot_out = b;
    return GS2D_OK;
}

int gs2d_frame_pipe_acquire(gs2d_frame_pipe* p, int slot, gs2d_stream_t stream, void** device_frame) {
    GS2D_REQUIRE(p && device_frame && slot >= 0 && slot < p->depth, "frame_pipe_acquire: bad argument");
    GS2D_CHECK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, p->copied[slot], 0));
    *device_frame = p->slots + (size_t)slot * p->frame_bytes;
    return GS2D_OK;
}

int gs2d_frame_pipe_release(gs2d_frame_pipe* p, int slot, gs2d_stream_t stream) {
    GS2D_REQUIRE(p && slot >= 0 && slot < p->depth, "frame_pipe_release: bad argument");
    GS2D_CHECK_CUDA(cudaEventRecord(p->freed[slot], (cudaStream_t)stream));
    p->has_consumer[slot] = true;
    return GS2D_OK;
}

}  // extern "C"
