// adan.cu -- the Adan step of the operator path as ONE kernel over all parameter tensors.
// Reference: optimizer.py:296-362 (_multi_tensor_adan: ~20 torch._foreach_* launches over the
// three or four parameter tensors of a frame model) -- and :365-395, where the reference itself
// anticipates an external fused kernel.  The fused fit step (fit.cu:k_fit_update) carries the same
// update inside its per-Gaussian kernel; this entry point serves callers that keep the autograd
// operator loop (gsvc_b200.adan.Adan(fused=True), FrameFit).
#include "gs2d_host.h"

namespace gs2d {

constexpr int kAdanMaxTensors = 8;
constexpr int kAdanThreads = 256;

struct AdanBatch {
    float* p[kAdanMaxTensors];
    const float* g[kAdanMaxTensors];
    float* m[kAdanMaxTensors];
    float* n[kAdanMaxTensors];
    float* d[kAdanMaxTensors];
    float* prev[kAdanMaxTensors];
    int first_block[kAdanMaxTensors + 1];   // blocks [first_block[t], first_block[t+1]) work on tensor t
    int count[kAdanMaxTensors];
    int num;
};

struct AdanScalars {
    float b1, b2, b3, step_size, step_size_diff, inv_bc3_sqrt, eps, decay_div;
    int first;   // previous gradient := gradient (optimizer.py:183-185)
    // step count in device memory (a captured CUDA graph replays the same arguments every iteration, so
    // the bias corrections cannot be baked into them): when set, the three step-dependent factors are
    // computed here from *step_dev with the arithmetic the host uses (double pow / sqrt, rounded once)
    const int* step_dev;
    float lr;
};

__device__ __forceinline__ void adan_bias_corrections(AdanScalars& h) {
    const double t = (double)*h.step_dev;
    h.step_size = (float)((double)h.lr / (1.0 - pow((double)h.b1, t)));
    h.step_size_diff = (float)((double)h.lr * (double)h.b2 / (1.0 - pow((double)h.b2, t)));
    h.inv_bc3_sqrt = (float)(1.0 / sqrt(1.0 - pow((double)h.b3, t)));
}

__global__ void __launch_bounds__(kAdanThreads)
k_adan_step(AdanBatch tb, AdanScalars h) {
    if (h.step_dev) {
        __shared__ float s_corr[3];
        if (threadIdx.x == 0) {
            adan_bias_corrections(h);
            s_corr[0] = h.step_size; s_corr[1] = h.step_size_diff; s_corr[2] = h.inv_bc3_sqrt;
        }
        __syncthreads();
        h.step_size = s_corr[0]; h.step_size_diff = s_corr[1]; h.inv_bc3_sqrt = s_corr[2];
    }
    int t = 0;
#pragma unroll
    for (int k = 1; k < kAdanMaxTensors; ++k)
        if (k < tb.num && (int)blockIdx.x >= tb.first_block[k]) t = k;
    // four consecutive elements per thread (16-byte accesses when the tensor allows it)
    const int base = ((int)blockIdx.x - tb.first_block[t]) * kAdanThreads * 4 + threadIdx.x * 4;
    const int cnt = tb.count[t];
    if (base >= cnt) return;
    float* p = tb.p[t] + base; const float* g = tb.g[t] + base;
    float* m = tb.m[t] + base; float* n = tb.n[t] + base; float* d = tb.d[t] + base; float* pv = tb.prev[t] + base;
    const bool vec = base + 4 <= cnt &&
                     ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)n | (uintptr_t)d | (uintptr_t)pv) & 15) == 0);
    float vp[4], vg[4], vm[4], vn[4], vd[4], vpv[4];
    const int k_end = min(4, cnt - base);
    if (vec) {
        *reinterpret_cast<float4*>(vp) = *reinterpret_cast<const float4*>(p);
        *reinterpret_cast<float4*>(vg) = *reinterpret_cast<const float4*>(g);
        *reinterpret_cast<float4*>(vm) = *reinterpret_cast<const float4*>(m);
        *reinterpret_cast<float4*>(vn) = *reinterpret_cast<const float4*>(n);
        *reinterpret_cast<float4*>(vd) = *reinterpret_cast<const float4*>(d);
        if (!h.first) *reinterpret_cast<float4*>(vpv) = *reinterpret_cast<const float4*>(pv);
    } else {
        for (int k = 0; k < k_end; ++k) {
            vp[k] = p[k]; vg[k] = g[k]; vm[k] = m[k]; vn[k] = n[k]; vd[k] = d[k];
            if (!h.first) vpv[k] = pv[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (k >= k_end) break;
        const float gr = vg[k];
        const float diff = h.first ? 0.f : gr - vpv[k];
        vm[k] = vm[k] * h.b1 + (1.f - h.b1) * gr;
        vd[k] = vd[k] * h.b2 + (1.f - h.b2) * diff;
        const float u = gr + h.b2 * diff;
        vn[k] = vn[k] * h.b3 + (1.f - h.b3) * u * u;
        const float denom = sqrtf(vn[k]) * h.inv_bc3_sqrt + h.eps;
        const float q = vp[k] - h.step_size * (vm[k] / denom) - h.step_size_diff * (vd[k] / denom);
        vp[k] = h.decay_div == 1.f ? q : q / h.decay_div;
    }
    if (vec) {
        *reinterpret_cast<float4*>(p) = *reinterpret_cast<float4*>(vp);
        *reinterpret_cast<float4*>(m) = *reinterpret_cast<float4*>(vm);
        *reinterpret_cast<float4*>(n) = *reinterpret_cast<float4*>(vn);
        *reinterpret_cast<float4*>(d) = *reinterpret_cast<float4*>(vd);
        *reinterpret_cast<float4*>(pv) = *reinterpret_cast<float4*>(vg);
    } else {
        for (int k = 0; k < k_end; ++k) { p[k] = vp[k]; m[k] = vm[k]; n[k] = vn[k]; d[k] = vd[k]; pv[k] = vg[k]; }
    }
}

}  // namespace gs2d

using namespace gs2d;

extern "C" int gs2d_adan_step(int num_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                              float* const* exp_avg_sq, float* const* exp_avg_diff, float* const* prev_grad,
                              const int* counts, float lr, float beta1, float beta2, float beta3, float eps,
                              float weight_decay, int step, const int* step_dev, int first, gs2d_stream_t stream) {
    GS2D_REQUIRE(num_tensors >= 0 && (step >= 1 || step_dev != nullptr), "adan_step: bad arguments");
    if (step < 1) step = 1;
    GS2D_REQUIRE(num_tensors == 0 || (params && grads && exp_avg && exp_avg_sq && exp_avg_diff && prev_grad && counts),
                 "adan_step: null pointer table");
    cudaStream_t s = (cudaStream_t)stream;
    AdanScalars h;
    h.b1 = beta1; h.b2 = beta2; h.b3 = beta3;
    h.step_size = (float)((double)lr / (1.0 - pow((double)beta1, step)));
    h.step_size_diff = (float)((double)lr * (double)beta2 / (1.0 - pow((double)beta2, step)));
    h.inv_bc3_sqrt = (float)(1.0 / sqrt(1.0 - pow((double)beta3, step)));
    h.eps = eps;
    h.decay_div = 1.f + lr * weight_decay;
    h.first = first != 0;
    h.step_dev = step_dev;
    h.lr = lr;
    for (int t0 = 0; t0 < num_tensors; t0 += kAdanMaxTensors) {
        AdanBatch tb = {};
        tb.num = num_tensors - t0 < kAdanMaxTensors ? num_tensors - t0 : kAdanMaxTensors;
        int blocks = 0;
        for (int k = 0; k < tb.num; ++k) {
            const int i = t0 + k;
            GS2D_REQUIRE(counts[i] >= 0 && (counts[i] == 0 || (params[i] && grads[i] && exp_avg[i] && exp_avg_sq[i] &&
                                                               exp_avg_diff[i] && prev_grad[i])),
                         "adan_step: null tensor");
            tb.p[k] = params[i]; tb.g[k] = grads[i]; tb.m[k] = exp_avg[i]; tb.n[k] = exp_avg_sq[i];
            tb.d[k] = exp_avg_diff[i]; tb.prev[k] = prev_grad[i];
            tb.count[k] = counts[i];
            tb.first_block[k] = blocks;
            blocks += div_up(counts[i], kAdanThreads * 4);
        }
        tb.first_block[tb.num] = blocks;
        if (blocks == 0) continue;
        ProfScope ps_(GS2D_STAGE_ADAN, s);
        k_adan_step<<<blocks, kAdanThreads, 0, s>>>(tb, h);
        GS2D_CHECK_LAUNCH("adan_step");
    }
    return GS2D_OK;
}
