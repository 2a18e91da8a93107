// pipe.cu -- frame pipe: raw frames from pinned host memory into a ring of device slots on a
// copy stream of its own, so that the copy of frame i+1 runs under the fit step of frame i.
// Pure host-side stream/event plumbing (six driver calls per frame); it exists natively because
// the same plumbing through a Python tensor library costs more host time per frame than the
// whole fit step takes on the device.
#include <cstdint>
#include <cstdlib>
#include <new>

#include "gs2d_host.h"

struct gs2d_frame_pipe {
    char* slots;
    size_t frame_bytes;
    int depth;
    int next;
    unsigned pushes;
    // One copy stream carries the frames in order.  (GS2D_PIPE_STREAMS=2 alternates between two, so that the
    // set-up of one frame's DMA runs under the other's transfer: measured again late in round 2,
    // profiles/r2_e2e_fill_probe.txt -- no faster in steady state, 62.8 vs 63.4 us per 1080p frame, and the first
    // frames of a stream then share the link and arrive later: 68.9 vs 73.5 us per step over a 20-frame stream.)
    cudaStream_t copy_streams[2];
    cudaEvent_t copied[GS2D_PIPE_MAX_DEPTH];
    cudaEvent_t freed[GS2D_PIPE_MAX_DEPTH];
    bool has_consumer[GS2D_PIPE_MAX_DEPTH];
    // optional: every frame is also converted into the tile kernel's own order (gs2d_fit_target_prepare)
    // on the copy stream, right behind its copy -- under the previous frame's fit step like the copy itself
    float* prepared;
    size_t prepared_bytes;
    int img_h, img_w, target_format;
    cudaStream_t prep_stream;                       // the conversion: a third stage of the pipe, behind the copy
    cudaEvent_t arrived[GS2D_PIPE_MAX_DEPTH];       // raw bytes of the slot are on the device
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
    p->pushes = 0;
    p->prepared = nullptr;
    p->prepared_bytes = 0;
    cudaError_t e = cudaStreamCreateWithFlags(&p->copy_streams[0], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->copy_streams[1], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->prep_stream, cudaStreamNonBlocking);
    for (int i = 0; i < depth && e == cudaSuccess; ++i) {
        p->has_consumer[i] = false;
        e = cudaEventCreateWithFlags(&p->arrived[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->copied[i], cudaEventDisableTiming);
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
    cudaStreamSynchronize(p->copy_streams[0]);
    cudaStreamSynchronize(p->copy_streams[1]);
    cudaStreamSynchronize(p->prep_stream);
    for (int i = 0; i < p->depth; ++i) {
        cudaEventDestroy(p->arrived[i]);
        cudaEventDestroy(p->copied[i]);
        cudaEventDestroy(p->freed[i]);
    }
    cudaStreamDestroy(p->copy_streams[0]);
    cudaStreamDestroy(p->copy_streams[1]);
    cudaStreamDestroy(p->prep_stream);
    delete p;
    return GS2D_OK;
}

int gs2d_frame_pipe_push(gs2d_frame_pipe* p, const void* host_frame, int* slot_out) {
    GS2D_REQUIRE(p && host_frame && slot_out, "frame_pipe_push: null argument");
    const int b = p->next;
    p->next = (b + 1) % p->depth;
    static const bool two_streams = [] { const char* e = getenv("GS2D_PIPE_STREAMS"); return e && e[0] == '2'; }();   // A/B runs
    cudaStream_t cs = p->copy_streams[two_streams ? (p->pushes++ & 1) : 0];
    if (p->has_consumer[b]) GS2D_CHECK_CUDA(cudaStreamWaitEvent(cs, p->freed[b], 0));
    GS2D_CHECK_CUDA(cudaMemcpyAsync(p->slots + (size_t)b * p->frame_bytes, host_frame, p->frame_bytes,
                                    cudaMemcpyHostToDevice, cs));
    if (p->prepared) {
        // third stage: the copy stream goes straight on to the next frame while this one is converted
        // (copy of frame i+2 | conversion of frame i+1 | fit step of frame i)
        GS2D_CHECK_CUDA(cudaEventRecord(p->arrived[b], cs));
        GS2D_CHECK_CUDA(cudaStreamWaitEvent(p->prep_stream, p->arrived[b], 0));
        const int rc = gs2d_fit_target_prepare(p->img_h, p->img_w, p->slots + (size_t)b * p->frame_bytes, p->target_format,
                                               (float*)((char*)p->prepared + (size_t)b * p->prepared_bytes),
                                               (gs2d_stream_t)p->prep_stream);
        if (rc != GS2D_OK) return rc;
        GS2D_CHECK_CUDA(cudaEventRecord(p->copied[b], p->prep_stream));
    } else {
        GS2D_CHECK_CUDA(cudaEventRecord(p->copied[b], cs));
    }
    *slot_out = b;
    return GS2D_OK;
}

int gs2d_frame_pipe_set_prepare(gs2d_frame_pipe* p, int img_h, int img_w, int target_format, float* prepared_slots) {
    GS2D_REQUIRE(p != nullptr, "frame_pipe_set_prepare: null pipe");
    if (prepared_slots == nullptr) { p->prepared = nullptr; return GS2D_OK; }
    GS2D_REQUIRE(img_h > 0 && img_w > 0 && ((uintptr_t)prepared_slots & 15) == 0, "frame_pipe_set_prepare: bad arguments");
    GS2D_REQUIRE(target_format == GS2D_TARGET_F32_CHW || target_format == GS2D_TARGET_U8_HWC ||
                     target_format == GS2D_TARGET_I420, "frame_pipe_set_prepare: unknown frame format");
    p->prepared = prepared_slots;
    p->prepared_bytes = gs2d_fit_target_bytes(img_h, img_w);
    p->img_h = img_h; p->img_w = img_w; p->target_format = target_format;
    return GS2D_OK;
}

int gs2d_frame_pipe_acquire(gs2d_frame_pipe* p, int slot, gs2d_stream_t stream, void** device_frame) {
    GS2D_REQUIRE(p && device_frame && slot >= 0 && slot < p->depth, "frame_pipe_acquire: bad argument");
    GS2D_CHECK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, p->copied[slot], 0));
    *device_frame = p->prepared ? (void*)((char*)p->prepared + (size_t)slot * p->prepared_bytes)
                                : (void*)(p->slots + (size_t)slot * p->frame_bytes);
    return GS2D_OK;
}

int gs2d_frame_pipe_release(gs2d_frame_pipe* p, int slot, gs2d_stream_t stream) {
    GS2D_REQUIRE(p && slot >= 0 && slot < p->depth, "frame_pipe_release: bad argument");
    GS2D_CHECK_CUDA(cudaEventRecord(p->freed[slot], (cudaStream_t)stream));
    p->has_consumer[slot] = true;
    return GS2D_OK;
}

}  // extern "C"
