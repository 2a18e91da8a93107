// density.cu -- device-side bookkeeping of the reference's prune control
// (GaussianSplats_Represent.py:98-128 removal_control, :130-172 adaptive_control): select the
// `remove_count` Gaussians with the smallest colour-weight norm and compact the parameter
// tensors, without a host round trip.  The reference does this with torch.norm + torch.sort +
// boolean-mask indexing of four tensors (about a dozen launches and a device->host sync for the
// mask's population count); here it is key build -> onesweep sort (this library's, 32 key bits)
// -> mark -> look-back scan (this library's) -> one gather kernel.
#include <cstdint>

#include "gs2d_host.h"

namespace gs2d {

constexpr int kThreads = 256;

// key = bits of ||rgb_w[i]||_2 over a row of ONE element: sqrt(w*w), each operation rounded to
// nearest as ATen's norm reduction does (no fast-math in torch) -- not |w|: the two differ in the
// last bit for some w, and the sort must see what the reference's sort sees.  Non-negative floats
// order like their bit patterns.
__global__ void __launch_bounds__(kThreads)
k_prune_keys(int n, const float* __restrict__ rgb_w, int64_t* __restrict__ keys, int32_t* __restrict__ vals,
             int32_t* __restrict__ keep) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const float w = rgb_w[i];
    const float nrm = __fsqrt_rn(__fmul_rn(w, w));
    keys[i] = (int64_t)__float_as_uint(nrm);
    vals[i] = i;
    keep[i] = 1;
}

__global__ void __launch_bounds__(kThreads)
k_prune_mark(int remove_count, const int32_t* __restrict__ order, int32_t* __restrict__ keep) {
    const int j = blockIdx.x * kThreads + threadIdx.x;
    if (j < remove_count) keep[order[j]] = 0;
}

__global__ void __launch_bounds__(kThreads)
k_prune_gather(int n, const int32_t* __restrict__ keep, const int32_t* __restrict__ cum,
               const float* __restrict__ rgb_w, const float* __restrict__ xyz, const float* __restrict__ chol,
               const float* __restrict__ feat, float* __restrict__ rgb_w_out, float* __restrict__ xyz_out,
               float* __restrict__ chol_out, float* __restrict__ feat_out, int32_t* __restrict__ keep_index_out) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= n || keep[i] == 0) return;
    const int d = cum[i] - 1;   // inclusive scan of the keep flags
    rgb_w_out[d] = rgb_w[i];
    xyz_out[2 * d] = xyz[2 * i]; xyz_out[2 * d + 1] = xyz[2 * i + 1];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        chol_out[3 * d + k] = chol[3 * i + k];
        feat_out[3 * d + k] = feat[3 * i + k];
    }
    if (keep_index_out) keep_index_out[d] = i;
}

static inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct PruneLayout {
    size_t keys_in, keys_out, vals_in, vals_out, keep, cum, sort, scan, total;
};

static PruneLayout prune_layout(int n) {
    PruneLayout L;
    const size_t m = (size_t)(n > 0 ? n : 1);
    size_t o = 0;
    L.keys_in = o;  o = up256(o + m * 8);
    L.keys_out = o; o = up256(o + m * 8);
    L.vals_in = o;  o = up256(o + m * 4);
    L.vals_out = o; o = up256(o + m * 4);
    L.keep = o;     o = up256(o + m * 4);
    L.cum = o;      o = up256(o + m * 4);
    L.sort = o;     o = up256(o + gs2d_sort_scratch_bytes(n));
    L.scan = o;     o = up256(o + gs2d_scan_scratch_bytes(n));
    L.total = o;
    return L;
}

}  // namespace gs2d

using namespace gs2d;

extern "C" {

size_t gs2d_prune_scratch_bytes(int n) { return prune_layout(n).total; }

int gs2d_prune_lowest(int n, int remove_count, const float* rgb_w, const float* xyz_raw, const float* chol_raw,
                      const float* features, float* rgb_w_out, float* xyz_out, float* chol_out, float* features_out,
                      int32_t* keep_index_out, void* scratch, size_t scratch_bytes, gs2d_stream_t stream) {
    GS2D_REQUIRE(n > 0 && remove_count >= 0 && remove_count <= n, "prune_lowest: bad counts");
    GS2D_REQUIRE(rgb_w && xyz_raw && chol_raw && features, "prune_lowest: null input");
    GS2D_REQUIRE(remove_count == n || (rgb_w_out && xyz_out && chol_out && features_out), "prune_lowest: null output");
    const PruneLayout L = prune_layout(n);
    GS2D_REQUIRE(scratch != nullptr && scratch_bytes >= L.total, "prune_lowest: scratch too small");
    GS2D_REQUIRE(((uintptr_t)scratch & 255) == 0, "prune_lowest: scratch must be 256-byte aligned");
    if (remove_count == n) return GS2D_OK;
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)scratch;
    int64_t* keys_in = (int64_t*)(w + L.keys_in);
    int64_t* keys_out = (int64_t*)(w + L.keys_out);
    int32_t* vals_in = (int32_t*)(w + L.vals_in);
    int32_t* vals_out = (int32_t*)(w + L.vals_out);
    int32_t* keep = (int32_t*)(w + L.keep);
    int32_t* cum = (int32_t*)(w + L.cum);
    const int blocks = div_up(n, kThreads);
    k_prune_keys<<<blocks, kThreads, 0, s>>>(n, rgb_w, keys_in, vals_in, keep);
    GS2D_CHECK_LAUNCH("prune_keys");
    if (remove_count > 0) {
        int rc = gs2d_sort_pairs(n, keys_in, vals_in, keys_out, vals_out, 0, 32, w + L.sort,
                                 gs2d_sort_scratch_bytes(n), stream);
        if (rc != GS2D_OK) return rc;
        k_prune_mark<<<div_up(remove_count, kThreads), kThreads, 0, s>>>(remove_count, vals_out, keep);
        GS2D_CHECK_LAUNCH("prune_mark");
    }
    int rc = gs2d_cumsum_tiles_hit(n, keep, cum, nullptr, nullptr, w + L.scan, gs2d_scan_scratch_bytes(n), stream);
    if (rc != GS2D_OK) return rc;
    k_prune_gather<<<blocks, kThreads, 0, s>>>(n, keep, cum, rgb_w, xyz_raw, chol_raw, features, rgb_w_out, xyz_out,
                                               chol_out, features_out, keep_index_out);
    GS2D_CHECK_LAUNCH("prune_gather");
    return GS2D_OK;
}

}  // extern "C"
