// fit.cu -- the fit iteration of one frame as TWO kernels (tile slots) or three (tile ranges), no host
// synchronisation, no allocation, no memset:
//
//   k_fit_tiles    one CTA per 16x16 tile: sum-rasterise forward (forward.cu:512-627, first 256
//                  entries only), clamp + L2 loss against the target and its
//                  gradient (GaussianSplats_Represent.py:88-89,194 + utils.py:21-26 of the
//                  reference) kept in shared memory, then the rasteriser backward
//                  (backward.cu:696-862) eight lanes per (Gaussian, tile) pair from shared memory:
//                  one 8-float reduction per pair.
//   k_fit_update   per Gaussian: projection backward (backward2d.cu:8-51, reference formula,
//                  SURVEY Q12), tanh backward, Adan update of the 8 (9 with rgb_W) parameters,
//                  re-zeroing of the gradient row, and -- with the updated parameters still in
//                  registers -- the NEXT iteration's binning: with tile slots it writes that
//                  iteration's packed records itself; with tile ranges it counts tiles and the last
//                  block scans the counts, and
//   k_fit_scatter  (ranges: every step; slots: only when nothing is pre-scattered) claims a position
//                  in each touched tile and writes the 48-byte packed record there.  Order inside a
//                  tile is restored in the tile kernel by ranking the records on Gaussian id, which
//                  is exactly the order the reference's stable sort on tile id produces.
//   k_fit_count    (ranges, first step only) per Gaussian: activations, projection, one atomic per
//                  touched tile; the last block to finish scans tile_count into tile_off (= the
//                  reference's tile_bins) and publishes num_intersects.
//
// The count/scatter/tile kernels without the loss and backward are the render-only path.
#include <cuda_fp16.h>
#include <type_traits>

#include "gs2d_internal.h"
#include "gs2d_math.cuh"
#include "gs2d_stage.cuh"

namespace gs2d {

#ifndef GS2D_FIT_THREADS
#define GS2D_FIT_THREADS 128   // 392 CTAs at 50k Gaussians: 2-3 per SM; 256 left 48 SMs with twice the work of the rest (update 20.9 -> 16.8 us)
#endif
constexpr int kFitThreads = GS2D_FIT_THREADS;   // per-Gaussian kernels (count, scatter, update)
constexpr int kStNumIntersects = 0;   // status words (int32[16])
constexpr int kStOverflow = 1;        // first step index whose intersections exceeded cap (0 = none)
constexpr int kStTicket = 2;
constexpr int kStMaxPerTile = 3;      // max entries of any tile in the last step
constexpr int kStSlotIsect = 4;       // [4], [5]: slot mode, intersections of the step with that parity
constexpr int kStSlotMax = 32;        // slot mode's running max per tile: one reduction per CTA, so it lives on
                                      // a cache line of its own (every CTA's first load is the flag in line 0)
constexpr int kScatterBatch = 8;      // slot claims a thread keeps in flight before it writes records
#ifndef GS2D_SLOT_STRIDE
#define GS2D_SLOT_STRIDE GS2D_MAX_PER_TILE
#endif
constexpr int kSlotStride = GS2D_SLOT_STRIDE;   // records between two tiles' slots (experiments only: < 256 drops the Q1 guarantee)

// Programmatic dependent launch (PDL): a kernel launched with the attribute may start while the
// kernel before it in the stream is still running; it must not touch what that kernel writes until
// pdl_wait() returns (= the predecessor finished and its writes are visible).  pdl_release() lets
// the NEXT kernel in the stream start early in turn; it is called after pdl_wait(), so a kernel's
// prologue never overlaps anything older than its direct predecessor.  Both are no-ops for a
// kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// (float)v / 255.f for v in 0..255, correctly rounded, without the division: q = v * fl(1/255),
// one FMA residual step (the recurrence is checked exhaustively against IEEE division in
// tests/test_capi_cpu.py; the device result in tests/test_gpu_fit.py)
__device__ __forceinline__ float u8_over_255(uint32_t v) {
    const float x = (float)v, r = 0.0039215688593685626984f;
    const float q = mulf(x, r);
    return fmaf_(fmaf_(-q, 255.f, x), r, q);
}

// cv2.COLOR_YUV2RGB_I420 for one pixel (utils.py:150 of the reference; OpenCV's BT.601 limited-range
// 20-bit fixed point; pinned against cv2 by tests/test_yuv_cpu.py) -> r | g << 8 | b << 16
__device__ __forceinline__ uint32_t yuv_to_rgb_packed(int y, int u, int v) {
    const int yy = max(y - 16, 0) * 1220542 + (1 << 19);
    const int uu = u - 128, vv = v - 128;
    const int r = (yy + 1673527 * vv) >> 20;
    const int g = (yy - 852492 * vv - 409993 * uu) >> 20;
    const int b = (yy + 2116026 * uu) >> 20;
    return (uint32_t)min(max(r, 0), 255) | ((uint32_t)min(max(g, 0), 255) << 8) | ((uint32_t)min(max(b, 0), 255) << 16);
}

// Where a Gaussian's parameters come from (`activate`):
//   0  means2d / L_elements / colours as the projection API takes them
//   1  the raw parameters of the frame model: tanh and the Cholesky bound applied here
//      (GaussianSplats_Represent.py:56-70)
//   2  the quantised wire format of the codec (forward_quantize, GaussianSplats_Compress.py:71-84):
//      xyz as fp16 bits (FakeQuantizationHalf, quantize.py:15-24), Cholesky as 6-bit codes with a
//      per-channel scale / beta (UniformQuantizer.compress / decompress, quantize.py:82-88:
//      code * scale + beta, then the bound), colours as two 3-bit residual-VQ indices into two
//      8-entry codebooks (VectorQuantizer.decompress, quantize.py:146-150: 0 + e0[i0] + e1[i1]).
//      The pointers then hold uint16 [n][2], uint8 [n][3] and uint8 [n][2]: 9 bytes per Gaussian
//      instead of 32, decoded in registers -- no fp32 copy of the frame's Gaussians ever exists.
//   3  the OUTPUTS of the projection operator (operator path: rasterize_gaussians_sum is handed
//      xys / radii / conics by its caller): nothing is projected again, the tile bbox comes from
//      (xys, radii) exactly as the reference's map_gaussian_to_intersects recomputes it
//      (forward.cu:113-123).  The pointers then hold xys [n][2], conics [n][3], colours [n][3] and
//      `radii` int32 [n].
struct QuantConst {
    float scale[3], beta[3];
    const float* codebooks;   // [2][8][3]
    const int32_t* radii;     // activate == 3 only
    int32_t* mail_host;       // operator path: pinned int32[8] the scan reports into (may be NULL)
    int mail_seq;
    // activate 0 / 1: floats between consecutive rows of the three parameter tensors; 0 = contiguous
    // (2 / 3 / 3).  8 / 8 / 8 with the pointers at +0 / +2 / +5 of one buffer = the packed [n][8] rows
    // of gs2d_fit_options.params8 (two aligned 16-byte loads per Gaussian)
    int ld_xyz, ld_chol, ld_feat;
};
static inline QuantConst packed8_rows() {
    QuantConst qc = QuantConst();
    qc.ld_xyz = qc.ld_chol = qc.ld_feat = 8;
    return qc;
}

__device__ __forceinline__ void fit_project(int i, const float* __restrict__ xyz_raw, const float* __restrict__ chol_raw,
                                            int activate, float half_w, float half_h, int tx, int ty, Proj2d& p,
                                            float2& means, float3& L, const QuantConst& qc) {
    if (activate == 3) {
        const float2 c = reinterpret_cast<const float2*>(xyz_raw)[i];
        const int rad = qc.radii[i];
        means = c;
        L = make_float3(0.f, 0.f, 0.f);
        p.cx = c.x; p.cy = c.y;
        p.conic = make_float3(chol_raw[3 * i], chol_raw[3 * i + 1], chol_raw[3 * i + 2]);
        p.radius = (float)rad;
        p.x0 = p.y0 = p.x1 = p.y1 = 0;
        p.nth = 0;
        p.ok = rad > 0;                   // forward.cu:111: radii <= 0 emits nothing
        if (p.ok) {
            tile_bbox(p.cx, p.cy, p.radius, tx, ty, p.x0, p.y0, p.x1, p.y1);
            const int area = (p.x1 - p.x0) * (p.y1 - p.y0);
            p.nth = area > 0 ? area : 0;
        }
        return;
    }
    if (activate == 2) {
        const uint32_t hx = reinterpret_cast<const uint32_t*>(xyz_raw)[i];          // two fp16
        const uint8_t* c = reinterpret_cast<const uint8_t*>(chol_raw) + 3 * (size_t)i;
        means = make_float2(tanhf(__half2float(__ushort_as_half((unsigned short)(hx & 0xffffu)))),
                            tanhf(__half2float(__ushort_as_half((unsigned short)(hx >> 16)))));
        L.x = addf(addf(mulf((float)c[0], qc.scale[0]), qc.beta[0]), 0.5f);
        L.y = addf(addf(mulf((float)c[1], qc.scale[1]), qc.beta[1]), 0.f);
        L.z = addf(addf(mulf((float)c[2], qc.scale[2]), qc.beta[2]), 0.5f);
    } else {
        float2 r;
        if (qc.ld_xyz == 8 && qc.ld_chol == 8 && chol_raw == xyz_raw + 2) {
            // packed rows: xyz(2) chol(3) feat(3) -- two 16-byte loads (the second one is shared with
            // fit_colour's: the compiler sees the same address)
            const float4 a = reinterpret_cast<const float4*>(xyz_raw)[2 * (size_t)i];
            const float4 b = reinterpret_cast<const float4*>(xyz_raw)[2 * (size_t)i + 1];
            r = make_float2(a.x, a.y);
            L = make_float3(a.z, a.w, b.x);
        } else {
            const size_t lx = qc.ld_xyz ? qc.ld_xyz : 2, lc = qc.ld_chol ? qc.ld_chol : 3;
            r = *reinterpret_cast<const float2*>(xyz_raw + lx * i);
            L = make_float3(chol_raw[lc * i], chol_raw[lc * i + 1], chol_raw[lc * i + 2]);
        }
        if (activate) {
            means = make_float2(tanhf(r.x), tanhf(r.y));
            L.x = addf(L.x, 0.5f); L.y = addf(L.y, 0.f); L.z = addf(L.z, 0.5f);
        } else {
            means = r;
        }
    }
    project_chol(means.x, means.y, L.x, L.y, L.z, half_w, half_h, tx, ty, p);
}

__device__ __forceinline__ float3 fit_colour(int i, const float* __restrict__ feat, const float* __restrict__ rgb_w,
                                             int activate, const QuantConst& qc) {
    if (activate == 2) {
        const uint8_t* ix = reinterpret_cast<const uint8_t*>(feat) + 2 * (size_t)i;
        const float* e0 = qc.codebooks + 3 * (ix[0] & 7), * e1 = qc.codebooks + 24 + 3 * (ix[1] & 7);
        return make_float3(addf(__ldg(e0), __ldg(e1)), addf(__ldg(e0 + 1), __ldg(e1 + 1)), addf(__ldg(e0 + 2), __ldg(e1 + 2)));
    }
    // colours = _features_dc * rgb_W (GaussianSplats_Represent.py:62-63); plain fp32 products
    const float cw = rgb_w ? rgb_w[i] : 1.f;
    if (qc.ld_feat == 8 && ((uintptr_t)(feat - 5) & 15) == 0) {   // packed rows: feat = row + 5
        const float4 b = reinterpret_cast<const float4*>(feat - 5)[2 * (size_t)i + 1];
        return make_float3(mulf(b.y, cw), mulf(b.z, cw), mulf(b.w, cw));
    }
    const size_t lf = qc.ld_feat ? qc.ld_feat : 3;
    return make_float3(mulf(feat[lf * i], cw), mulf(feat[lf * i + 1], cw), mulf(feat[lf * i + 2], cw));
}

// ------------------------------------------------------------------------------------------
// K1: count + (last block) scan of the per-tile counts
// ------------------------------------------------------------------------------------------
// [row0, row1): the band of tile rows this rank rasterises (the whole frame unless the frame is
// split across GPUs by tile rows, SURVEY 8e spatial mode)
// (first, stride): the tiles of the Gaussian's bbox are numbered row by row; this call takes numbers first,
// first + stride, ... -- several lanes can share one Gaussian's tiles
// j / bw for the tile numbering of a bbox (0 <= j, 1 <= bw, both far below 2^23): one multiply by the
// reciprocal and a one-step correction instead of the ~20 instructions of an integer division by a variable;
// exact whatever the rounding of the reciprocal (the quotient estimate is off by at most one).
__device__ __forceinline__ int div_small(int j, int bw, float inv_bw, int& rem) {
    int dy = (int)((float)j * inv_bw);
    rem = j - dy * bw;
    if (rem < 0) { --dy; rem += bw; }
    else if (rem >= bw) { ++dy; rem -= bw; }
    return dy;
}

__device__ __forceinline__ void count_tiles(const Proj2d& p, int tx, int row0, int row1, int32_t* tile_count,
                                            int first = 0, int stride = 1) {
    if (p.nth <= 0) return;
    const int ya = max(p.y0, row0), yb = min(p.y1, row1), bw = p.x1 - p.x0;
    const int nt = (yb - ya) * bw;
    const float inv_bw = __fdividef(1.f, (float)bw);
    for (int j = first; j < nt; j += stride) {
        int dx;
        const int dy = div_small(j, bw, inv_bw, dx);
        atomicAdd(&tile_count[(ya + dy) * tx + p.x0 + dx], 1);
    }
}

// Called by every thread of every block after its count_tiles(): the last block to arrive turns
// the per-tile counts into tile ranges.  sh: 8 ints, s_last: 1 int of shared memory.
constexpr int kScanPerThread = 64;   // tiles a thread of the scanning block owns per pass
__device__ __forceinline__ void scan_tiles_by_last_block(int tx, int ty, int cap, int step, int32_t* tile_count,
                                                         int32_t* __restrict__ tile_off, int32_t* status, int* sh,
                                                         int* s_last, int32_t* mail_host = nullptr, int mail_seq = 0) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *s_last = (atomicAdd(&status[kStTicket], 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!*s_last) return;
    __threadfence();
    // exclusive scan of tile_count over all tiles by this one block; counts are re-zeroed so the
    // scatter kernel can use the same array as its per-tile fill counters.  Each thread owns 64
    // consecutive tiles per pass and keeps them in registers: sixteen 16-byte loads in flight at
    // once -- the 8 160 tiles of a 1080p frame are ONE pass, i.e. one L2 round trip, of this one block.
    const int tiles = tx * ty;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0, mx = 0;
    for (int base = 0; base < tiles; base += kFitThreads * kScanPerThread) {
        const int t0 = base + threadIdx.x * kScanPerThread;
        int v[kScanPerThread];
#pragma unroll
        for (int k = 0; k < kScanPerThread / 4; ++k) {
            const int t = t0 + 4 * k;
            if (t + 3 < tiles) {
                const int4 q = __ldcg(reinterpret_cast<const int4*>(tile_count + t));
                v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) v[4 * k + j] = (t + j < tiles) ? __ldcg(tile_count + t + j) : 0;
            }
        }
        int sum = 0;
#pragma unroll
        for (int j = 0; j < kScanPerThread; ++j) {
            const int c = v[j];
            mx = max(mx, c);
            v[j] = sum;      // exclusive within the thread
            sum += c;
        }
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        __syncthreads();   // sh is reused across passes
        if (lane == 31) sh[warp] = incl;
        __syncthreads();
        int wexcl = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kFitThreads / 32; ++w) {
            const int x = sh[w];
            if (w < warp) wexcl += x;
            total += x;
        }
        const int excl = carry + wexcl + incl - sum;
        const int4 z = make_int4(0, 0, 0, 0);
#pragma unroll
        for (int k = 0; k < kScanPerThread / 4; ++k) {
            const int t = t0 + 4 * k;
            if (t + 3 < tiles) {
                *reinterpret_cast<int4*>(tile_off + t) =
                    make_int4(excl + v[4 * k], excl + v[4 * k + 1], excl + v[4 * k + 2], excl + v[4 * k + 3]);
                *reinterpret_cast<int4*>(tile_count + t) = z;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (t + j < tiles) { tile_off[t + j] = excl + v[4 * k + j]; tile_count[t + j] = 0; }
            }
        }
        carry += total;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __syncthreads();
    if (lane == 0) sh[warp] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kFitThreads / 32; ++w) mx = max(mx, sh[w]);
        tile_off[tiles] = carry;
        status[kStNumIntersects] = carry;
        status[kStMaxPerTile] = mx;
        status[kStTicket] = 0;
        if (carry > cap && status[kStOverflow] == 0) status[kStOverflow] = step > 0 ? step : 1;
        if (mail_host) {
            // operator path: the host learns the intersection count (capacity hint of later calls,
            // overflow check) from pinned memory whenever it next looks -- nothing waits for it
            mail_host[0] = carry;
            mail_host[1] = carry > cap ? 1 : 0;
            mail_host[2] = mx;
            __threadfence_system();
            mail_host[4] = mail_seq;      // written last: the record is complete once it shows
        }
    }
}

__global__ void __launch_bounds__(kFitThreads)
k_fit_count(int n, const float* __restrict__ xyz_raw, const float* __restrict__ chol_raw, int activate, float half_w,
            float half_h, int tx, int ty, int row0, int row1, int cap, int step, int32_t* tile_count,
            int32_t* __restrict__ tile_off, int32_t* status, QuantConst qc) {
    __shared__ int sh[8];
    __shared__ int s_last;
    const int i = blockIdx.x * kFitThreads + threadIdx.x;
    pdl_release();   // render chain: the scatter kernel's projection may run under this kernel (it waits before it reads the scan)
    if (i < n) {
        Proj2d p; float2 m; float3 L;
        fit_project(i, xyz_raw, chol_raw, activate, half_w, half_h, tx, ty, p, m, L, qc);
        count_tiles(p, tx, row0, row1, tile_count);
    }
    scan_tiles_by_last_block(tx, ty, cap, step, tile_count, tile_off, status, sh, &s_last, qc.mail_host, qc.mail_seq);
}

// ------------------------------------------------------------------------------------------
// K2: scatter the packed records into their tiles
// packed[pos] = { x, y, conic.x, conic.y | conic.z, opacity, rect, gaussian id | r, g, b, tile }
// Two addressings of the same records:
//   ranges (SLOT = false)  tile t owns [tile_off[t], tile_off[t+1]) -- needs the count + scan first;
//   slots  (SLOT = true)   tile t owns [t*256, t*256 + 256): no count, no scan, so the update kernel
//                          can scatter the NEXT step's records itself (two kernels per iteration).
//                          256 is all the reference ever renders of a tile; WHICH 256
//                          of a fuller tile needs all its ids, so a 257th claim flags the step
//                          (status[kStOverflow] = -step) and the caller redoes it with ranges.
// The claims (atomics that return a position) of up to kScatterBatch tiles are issued back to back
// before the first record is written: a Gaussian's tiles then cost one round trip to L2, not one each.
// Returns the number of tiles claimed.
// ------------------------------------------------------------------------------------------
// Sharded update (include/gs2d.h, gs2d_grad_exchange): the peers' parameter tensors (the update
// kernel stores its shard's new values into all of them) and status words (an overflowing tile is
// flagged on every rank); R == 0: not sharded.
struct Route {
    float* const* xyz;
    float* const* chol;
    float* const* feat;
    int32_t* const* status;
    int R;
    // remote scatter (gs2d_grad_exchange.work_peers): the ranks' workspaces, all of one layout, and the
    // tile-row edges of their bands -- the update kernel writes the NEXT step's records of every Gaussian
    // of its shard straight into the slots of the rank that rasterises the tile (NVLink atomics + stores)
    char* const* work;
    size_t off_tile_count, off_packed;
    const int32_t* band_edges;   // [R + 1]
};

// Slot mode's per-tile counter packs two counts into its 32 bits: bits 16.. = all (Gaussian, tile) pairs
// claimed for the tile -- the reference's count: a 257th flags the step, and it is what max_per_tile
// reports -- and bits 0..15 = the records actually STORED.  With `cull` the pairs whose pixel rectangle
// inside the tile is empty (pair_rect: no pixel of the tile can pass the alpha test, so the pair adds
// nothing to the image and gets no gradient; 23 % of the pairs at 1080p / 50k as initialised) are counted
// but not stored: the tile kernel stages, ranks and walks a quarter fewer records.  While every tile has at
// most 256 pairs all of them are rendered, so dropping the empty ones cannot change which are -- same results.
constexpr uint32_t kFillAll = 0x10000u;      // one more pair
constexpr uint32_t kFillStored = 0x1u;       // ... that also takes a slot
__device__ __forceinline__ bool rect_is_empty(uint32_t rect) { return (rect >> 24) < ((rect >> 16) & 255u); }

template <bool SLOT, int BATCH = kScatterBatch>
__device__ __forceinline__ int scatter_records(const Proj2d& p, int i, float o, float r, float g, float b, int tx,
                                               int row0, int row1, int W, int H, const int32_t* __restrict__ tile_off,
                                               int32_t* tile_fill, float4* __restrict__ packed, int32_t* status,
                                               int flag_step, const Route& rt = Route(), bool cull = false,
                                               int first = 0, int stride = 1) {
    // (first, stride): as count_tiles -- this call handles the bbox's tiles number first, first + stride, ...
    if (p.nth <= 0) return 0;
    const int ya = max(p.y0, row0), yb = min(p.y1, row1);
    const int bw = p.x1 - p.x0;
    const int nt = (yb - ya) * bw;
    if (nt <= first) return 0;
    int mine = 0;
    const PairExtent ext = pair_extent(p.cx, p.cy, o, p.conic.x, p.conic.y, p.conic.z);   // once per Gaussian, not per tile
    const float inv_bw = __fdividef(1.f, (float)bw);
    for (int base = first; base < nt; base += BATCH * stride) {
        int pos[BATCH], tid_[BATCH];
        uint32_t rect[BATCH];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            const int j = base + k * stride;
            if (j < nt) {
                int dx;
                const int dy = div_small(j, bw, inv_bw, dx);
                const int x = p.x0 + dx, y = ya + dy;
                tid_[k] = y * tx + x;
                ++mine;
                rect[k] = pair_rect_of(p.cx, p.cy, ext, x * GS2D_TILE, y * GS2D_TILE, W, H);
                if (SLOT) {
                    const bool store = !(cull && rect_is_empty(rect[k]));
                    const uint32_t old = atomicAdd(reinterpret_cast<uint32_t*>(&tile_fill[tid_[k]]),
                                                   store ? kFillAll + kFillStored : kFillAll);
                    // (all pairs so far) in the high half decides the overflow; a dropped pair gets position -1
                    pos[k] = (old >> 16) >= (uint32_t)kSlotStride ? kSlotStride : (store ? (int)(old & 0xffffu) : -1);
                } else {
                    pos[k] = atomicAdd(&tile_fill[tid_[k]], 1);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            if (base + k * stride < nt) {
                const int t = tid_[k];
                size_t at;
                bool ok = true;
                if (SLOT) {
                    ok = pos[k] >= 0 && pos[k] < kSlotStride;
                    if (pos[k] >= kSlotStride) {
                        atomicCAS(&status[kStOverflow], 0, -flag_step);
                        for (int q = 0; q < rt.R; ++q)   // sharded update: every rank must stop at this step
                            if (rt.status[q] != status) atomicCAS_system(&rt.status[q][kStOverflow], 0, -flag_step);
                    }
                    at = (size_t)t * kSlotStride + (ok ? pos[k] : 0);
                } else {
                    at = (size_t)tile_off[t] + pos[k];
                    GS2D_ASSERT(pos[k] >= 0 && at < (size_t)tile_off[t + 1]);   // the count kernel saw the same projection
                }
                if (ok) {
                    // the diagonal conic terms are stored HALVED: sigma = 0.5*(cx dx^2 + cz dy^2) + cy dx dy is
                    // then fma(dy, cy*dx, fma(dx, (cx/2)*dx, ((cz/2)*dy)*dy)) with no separate multiply by 0.5 per
                    // pixel; halving is exact, so every rounding is the reference's
                    packed[3 * at + 0] = make_float4(p.cx, p.cy, 0.5f * p.conic.x, p.conic.y);
                    packed[3 * at + 1] = make_float4(0.5f * p.conic.z, o, __uint_as_float(rect[k]), __int_as_float(i));
                    packed[3 * at + 2] = make_float4(r, g, b, __uint_as_float((uint32_t)t));
                }
            }
        }
    }
    return mine;
}

// scatter_records<true> across ranks: tile (x, y) belongs to the rank whose band holds row y; its fill
// counter and its slots are addressed in that rank's workspace (peer-mapped memory).  The claims are
// system-scope atomics over NVLink, up to kScatterBatch of them in flight per thread, the records plain
// stores that the cross-rank barrier after the update kernel publishes.  Returns the tiles claimed.
__device__ __forceinline__ int scatter_records_remote(const Proj2d& p, int i, float o, float r, float g, float b, int tx,
                                                      int W, int H, int32_t* status, int flag_step, const Route& rt) {
    if (p.nth <= 0) return 0;
    const int bw = p.x1 - p.x0;
    const int nt = (p.y1 - p.y0) * bw;
    if (nt <= 0) return 0;
    const PairExtent ext = pair_extent(p.cx, p.cy, o, p.conic.x, p.conic.y, p.conic.z);
    int x = p.x0, y = p.y0, q = 0;
    for (int base = 0; base < nt; base += kScatterBatch) {
        int pos[kScatterBatch];
        uint32_t rect[kScatterBatch];
        const int xs = x, ys = y, qs = q;
#pragma unroll
        for (int k = 0; k < kScatterBatch; ++k) {
            if (base + k < nt) {
                while (y >= rt.band_edges[q + 1]) ++q;     // rows only grow: the owner is found incrementally
                uint32_t* fill = reinterpret_cast<uint32_t*>(rt.work[q] + rt.off_tile_count);
                rect[k] = pair_rect_of(p.cx, p.cy, ext, x * GS2D_TILE, y * GS2D_TILE, W, H);
                const bool store = !rect_is_empty(rect[k]);
                const uint32_t old = atomicAdd_system(&fill[y * tx + x], store ? kFillAll + kFillStored : kFillAll);
                pos[k] = (old >> 16) >= (uint32_t)kSlotStride ? kSlotStride : (store ? (int)(old & 0xffffu) : -1);
                if (++x == p.x1) { x = p.x0; ++y; }
            }
        }
        int xr = xs, yr = ys, qr = qs;
#pragma unroll
        for (int k = 0; k < kScatterBatch; ++k) {
            if (base + k < nt) {
                while (yr >= rt.band_edges[qr + 1]) ++qr;
                const int t = yr * tx + xr;
                if (pos[k] < 0) {
                    // empty rectangle: counted, not stored
                } else if (pos[k] < kSlotStride) {
                    float4* packed = reinterpret_cast<float4*>(rt.work[qr] + rt.off_packed);
                    const size_t at = (size_t)t * kSlotStride + pos[k];
                    packed[3 * at + 0] = make_float4(p.cx, p.cy, 0.5f * p.conic.x, p.conic.y);
                    packed[3 * at + 1] = make_float4(0.5f * p.conic.z, o, __uint_as_float(rect[k]), __int_as_float(i));
                    packed[3 * at + 2] = make_float4(r, g, b, __uint_as_float((uint32_t)t));
                } else {
                    atomicCAS(&status[kStOverflow], 0, -flag_step);
                    for (int z = 0; z < rt.R; ++z)   // every rank must stop at this step
                        if (rt.status[z] != status) atomicCAS_system(&rt.status[z][kStOverflow], 0, -flag_step);
                }
                if (++xr == p.x1) { xr = p.x0; ++yr; }
            }
        }
    }
    return nt;
}

// block sum of one int per thread -> one atomic per block (slot mode's intersection count)
__device__ __forceinline__ void block_add(int v, int32_t* dst, int* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
        for (int w = 0; w < kFitThreads / 32; ++w) s += sh[w];
        if (s) atomicAdd(dst, s);
    }
}

// MULTI = false: no cross-rank overflow signalling compiled in (single GPU: the common case; these
// per-Gaussian kernels run every instruction once per warp, so code size is time)
template <bool MULTI>
__global__ void __launch_bounds__(kFitThreads)
k_fit_scatter(int n, const float* __restrict__ xyz_raw, const float* __restrict__ chol_raw,
              const float* __restrict__ feat, const float* __restrict__ opac, const float* __restrict__ rgb_w,
              int activate, float half_w, float half_h,
              int tx, int ty, int row0, int row1, int W, int H, const int32_t* __restrict__ tile_off,
              int32_t* tile_fill, int32_t* status, float4* __restrict__ packed, int slots, int step, QuantConst qc,
              Route rt) {
    __shared__ int sh[8];
    // not an early return: in slot mode another block may raise the flag while this one runs, and
    // every thread must reach the block sum below
    const int i = blockIdx.x * kFitThreads + threadIdx.x;
    int claimed = 0;
    // the projection depends on the parameters alone: launched programmatically after the count kernel
    // (render chain) it runs under that kernel; everything the count / scan wrote is read after the wait
    Proj2d p; float2 m; float3 L;
    float o = 1.f;
    float3 col = make_float3(0.f, 0.f, 0.f);
    // (MULTI, the sharded update of the tile-row split: the parameters are what the PEERS' update kernels stored
    // into this rank's tensors before the barrier kernel this launch depends on -- nothing is read before the wait)
    if (!MULTI && i < n) {
        fit_project(i, xyz_raw, chol_raw, activate, half_w, half_h, tx, ty, p, m, L, qc);
        o = opac ? opac[i] : 1.f;
        col = fit_colour(i, feat, rgb_w, activate, qc);
    }
    pdl_wait();
    pdl_release();
    if (MULTI && i < n) {
        fit_project(i, xyz_raw, chol_raw, activate, half_w, half_h, tx, ty, p, m, L, qc);
        o = opac ? opac[i] : 1.f;
        col = fit_colour(i, feat, rgb_w, activate, qc);
    }
    const bool dead = status[kStOverflow] != 0;
    if (i < n && !dead) {
        const float r = col.x, g = col.y, b = col.z;
        if (slots)
            claimed = scatter_records<true, 4>(p, i, o, r, g, b, tx, row0, row1, W, H, nullptr, tile_fill, packed, status, step,
                                               MULTI ? rt : Route(), (slots & 2) != 0);
        else
            scatter_records<false, 4>(p, i, o, r, g, b, tx, row0, row1, W, H, tile_off, tile_fill, packed, status, step);
    }
    if (slots) block_add(claimed, &status[kStSlotIsect + (step & 1)], sh);
}

// slot mode, parameters changed behind the library's back: forget the stale scatter
__global__ void __launch_bounds__(kFitThreads)
k_fit_reset_slots(int tiles, int32_t* tile_fill, int32_t* status) {
    const int t = blockIdx.x * kFitThreads + threadIdx.x;
    if (t < tiles) tile_fill[t] = 0;
    if (t == 0) { status[kStSlotIsect] = 0; status[kStSlotIsect + 1] = 0; status[kStSlotMax] = 0; }
}

// ------------------------------------------------------------------------------------------
// K3: the tile kernel
// ------------------------------------------------------------------------------------------
#ifndef GS2D_TILE_THREADS
#define GS2D_TILE_THREADS 64
#endif
constexpr int kTileThreads = GS2D_TILE_THREADS;            // warps per 16x16 tile x 32
constexpr int kPix = GS2D_TILE_PIX / kTileThreads;          // pixels per thread (one column, rows 2 apart)
constexpr int kWarpRows = 2 * kPix;                         // tile rows owned by one warp
static_assert(kTileThreads == 32 || kTileThreads == 64, "one or two warps per tile");
#ifndef GS2D_BWD_ROWS
#define GS2D_BWD_ROWS 2
#endif
#ifndef GS2D_BWD_UNROLL
#define GS2D_BWD_UNROLL 1   // trips unrolled by the compiler.  Measured (fit_tiles, C2 / trained-like): 1 -> 58.9 / 158.8 us at 70
                            // registers, 2 -> 59.7 / 160.2 at 80, left to the compiler (4) -> 65.2 / 166.3 at 103
#endif
constexpr int kBwdUnroll = GS2D_BWD_UNROLL;
constexpr int kBwdRows = GS2D_BWD_ROWS;   // rows of a pair's rectangle a lane evaluates per loop trip
#ifndef GS2D_WARP_STAGE
#define GS2D_WARP_STAGE 1   // tiles of at most 32 records (the common case: mean 15, max 31 at 1080p / 50k) are staged and
                            // ranked by each warp for itself -- registers + shuffles, no block barrier (stage_tile_warp)
#endif
#ifndef GS2D_TILE_PERSIST
#define GS2D_TILE_PERSIST 0   // experiment: see k_fit_tiles
#endif
#ifndef GS2D_TILE_PERSIST_OCC
#define GS2D_TILE_PERSIST_OCC 10   // resident CTAs per SM the persistent grid is sized for
#endif
#ifndef GS2D_SQ_RECORDS
#define GS2D_SQ_RECORDS GS2D_MAX_PER_TILE   // experiments only: a smaller batch buffer (tiles beyond it are then truncated)
#endif
constexpr int kWarpStage = GS2D_WARP_STAGE;
// experiment (VERDICT r1 #3e; profiles/r2_ab_bulk_stage.txt): slot mode, the tile's first 32 records arrive by ONE
// bulk asynchronous copy (cp.async.bulk, the 1-D form of TMA) into shared memory, completion on an mbarrier; the
// warps then rank out of shared memory.  Measured against the lanes' own 16-byte loads: see the file.  Off.
// gradient rows added by 16-byte vector reductions (REDG.E.ADD.F32x4): 0 never, 1 in the pushed-reduction
// instantiation (the additions cross NVLink), 2 always.  Measured: profiles/r2_ab_red_v4.txt
#ifndef GS2D_RED_V4
#define GS2D_RED_V4 1
#endif
constexpr int kRedV4 = GS2D_RED_V4;
#ifndef GS2D_STAGE_BULK
#define GS2D_STAGE_BULK 0
#endif
#if GS2D_STAGE_BULK
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load_expect(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar), d = (uint32_t)__cvta_generic_to_shared(dst);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(d), "l"(src), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(b), "r"(parity) : "memory");
}
#endif
constexpr int kSqRecords = GS2D_SQ_RECORDS > 32 * (kTileThreads / 32) ? GS2D_SQ_RECORDS : 32 * (kTileThreads / 32);

// The target frame in the tile kernel's own order: fp32 [tiles][kTileThreads][kPix][3], i.e. for every
// tile the 12 (kPix * 3) floats each thread compares its pixels with, contiguous.  A frame is fitted for
// thousands of iterations (train_video_Represent.py:85-86 of the reference); decoding its bytes in every
// one of them was a fifth of the tile kernel's instructions.  Values are what the kernel's own decode
// produces (same device functions), pixels outside the image are zero.
template <int GT>
__global__ void __launch_bounds__(kTileThreads)
k_target_tiled(int W, int H, int tx, const void* __restrict__ gt, float* __restrict__ out) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = blockIdx.y * tx + blockIdx.x;
    const int lx = lane & 15, ly0 = kWarpRows * warp + (lane >> 4);
    const int px_i = blockIdx.x * GS2D_TILE + lx;
    float tv[kPix * 3];
#pragma unroll
    for (int j = 0; j < kPix; ++j) {
        const int py_i = blockIdx.y * GS2D_TILE + ly0 + 2 * j;
        float3 v = make_float3(0.f, 0.f, 0.f);
        if (px_i < W && py_i < H) {
            const size_t pid = (size_t)py_i * W + px_i, plane = (size_t)H * W;
            if (GT == 0) {
                const float* g = (const float*)gt;
                v = make_float3(__ldg(g + pid), __ldg(g + plane + pid), __ldg(g + 2 * plane + pid));
            } else {
                uint32_t u;
                if (GT == 1) {
                    const uint8_t* g = (const uint8_t*)gt + 3 * pid;
                    u = (uint32_t)__ldg(g) | ((uint32_t)__ldg(g + 1) << 8) | ((uint32_t)__ldg(g + 2) << 16);
                } else {
                    const uint8_t* g = (const uint8_t*)gt;
                    const size_t cid = (size_t)(py_i >> 1) * (W >> 1) + (px_i >> 1);
                    u = yuv_to_rgb_packed(__ldg(g + pid), __ldg(g + plane + cid), __ldg(g + plane + (plane >> 2) + cid));
                }
                v = make_float3(u8_over_255(u & 255u), u8_over_255((u >> 8) & 255u), u8_over_255((u >> 16) & 255u));
            }
        }
        tv[3 * j] = v.x; tv[3 * j + 1] = v.y; tv[3 * j + 2] = v.z;
    }
    float4* o = reinterpret_cast<float4*>(out) + ((size_t)t * kTileThreads + tid) * (kPix * 3 / 4);
#pragma unroll
    for (int q = 0; q < kPix * 3 / 4; ++q) o[q] = *reinterpret_cast<float4*>(tv + 4 * q);
}

// MODE 0: fit (loss + backward, no image written unless out_img != NULL)
// MODE 1: render only (image [+ final_idx] written, no target, no backward)
// GT: 0 = target is fp32 planar [3][H][W] (what torchvision's ToTensor / the reference's loader
//         yields), 1 = uint8 interleaved [H][W][3] (raw frame; compared as value / 255),
//         2 = planar I420 as the .yuv file holds it (converted here as the reference's loader does)
//         3 = fp32 in this kernel's own per-thread order (gs2d_fit_target_prepare: any of the above,
//             converted once per frame with the same arithmetic -- same bits, no per-iteration decode)
//
// One CTA of two warps per tile.  Warp w owns tile rows 8w .. 8w+7; a lane owns one column of
// two interleaved row sets: pixel j (0..3) of lane l is (x = l & 15, y = 8w + 2j + (l >> 4)), so
// one warp instruction on pixel j covers the row pair {8w+2j, 8w+2j+1} -- the granularity of the
// rectangle culling -- while dx, conic.x*dx and conic.y*dx are computed once per Gaussian and
// reused by the four pixels (same operations as the reference's expression, so bit-identical).
#ifndef GS2D_TILE_MIN_BLOCKS
#define GS2D_TILE_MIN_BLOCKS 1
#endif
constexpr int kTileAllPerTile = 2;    //        every Gaussian of a tile is rendered, in batches of 256 (ranges only)
// ALL: the instantiation that walks a fuller tile in batches (kTileAllPerTile); a run-time switch in the one
// kernel cost the common case 42 registers and 12 % (58.9 -> 66.2 us), so it is a kernel of its own.
// PUSH: the instantiation of the pushed reduce-scatter (push_peers; the single-GPU kernel keeps its code).
template <int MODE, int GT, bool ALL = false, bool PUSH = false>
__global__ void __launch_bounds__(kTileThreads, GS2D_TILE_MIN_BLOCKS)
k_fit_tiles(int W, int H, int tx, int row0, const int32_t* __restrict__ tile_off, int32_t* tile_fill,
            const float4* __restrict__ packed, const void* __restrict__ gt, float grad_scale,
            float* __restrict__ grads, double* loss_acc, int32_t* __restrict__ status,
            float* __restrict__ out_img, int out_chw, int out_clamp, int32_t* __restrict__ final_idx,
            int32_t* __restrict__ ids_sorted_dbg, int step, int flags, float* const* __restrict__ push_peers,
            int push_per, float* __restrict__ loss_pad) {
    __shared__ float4 sq_all[kSqRecords * 3];
    // v_out of the tile (rows 0..15; before that, the id chunk of the staging) + 16 rows of zeros:
    // the backward's row walk may run past the tile for lanes whose pair has a shorter rectangle
    __shared__ float4 sv[2 * GS2D_TILE_PIX];
    __shared__ float sred[kTileThreads / 32];
    int* sid = reinterpret_cast<int*>(sv);
    static_assert(sizeof(float4) * GS2D_TILE_PIX >= sizeof(int) * kIdChunk, "id chunk must fit in sv");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#if GS2D_TILE_PERSIST
    // experiment: resident CTAs walk the tiles with a grid stride (flags >> 8 = tile rows of the band)
    const int n_tiles = tx * (flags >> 8);
    for (int tt = blockIdx.x; tt < n_tiles; tt += gridDim.x) {
    const int bx = tt % tx, by = tt / tx;
    if (tt != (int)blockIdx.x) __syncthreads();   // the previous tile's sv is dead
#define GS2D_TILE_DONE continue
#else
    const int bx = blockIdx.x, by = blockIdx.y;
#define GS2D_TILE_DONE return
#endif
    const int tile_y = by + row0;
    const int t = tile_y * tx + bx;
    if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < GS2D_TILE_PIX / kTileThreads; ++j)
            sv[GS2D_TILE_PIX + j * kTileThreads + tid] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#if GS2D_STAGE_BULK
    __shared__ __align__(128) float4 sraw[32 * 3];
    __shared__ __align__(8) unsigned long long sbar;
    if (tid == 0) mbar_init(&sbar, 1);
    __syncthreads();   // (in the prologue: before the wait for the previous kernel)
#endif

    const int lx = lane & 15, ly0 = kWarpRows * warp + (lane >> 4);   // pixel j: (lx, ly0 + 2j)
    const int ox = bx * GS2D_TILE, oy = tile_y * GS2D_TILE;
    const int px_i = ox + lx;
    const bool in_x = px_i < W;
    uint32_t tgt_u8[kPix] = {};
    uint8_t raw_y[kPix] = {}, raw_u[kPix] = {}, raw_v[kPix] = {};   // I420: decoded where the target is used
    float3 tgt[kPix];
#pragma unroll
    for (int j = 0; j < kPix; ++j) tgt[j] = make_float3(0.f, 0.f, 0.f);
    if (MODE == 0 && GT == 3) {
        // the frame as gs2d_fit_target_prepare laid it out: this thread's kPix pixels' (r, g, b) as
        // consecutive floats -- 16-byte loads, a warp reads one contiguous run, nothing to decode
        const float4* tp = reinterpret_cast<const float4*>(gt) + ((size_t)t * kTileThreads + tid) * (kPix * 3 / 4);
        float tv[kPix * 3];
#pragma unroll
        for (int q = 0; q < kPix * 3 / 4; ++q) *reinterpret_cast<float4*>(tv + 4 * q) = __ldg(tp + q);
#pragma unroll
        for (int j = 0; j < kPix; ++j) tgt[j] = make_float3(tv[3 * j], tv[3 * j + 1], tv[3 * j + 2]);
    }
    if (MODE == 0 && GT != 3 && in_x) {   // issued before the batch arrives: independent loads
#pragma unroll
        for (int j = 0; j < kPix; ++j) {
            const int py_i = oy + ly0 + 2 * j;
            if (py_i < H) {
                const size_t pid = (size_t)py_i * W + px_i;
                if (GT == 0) {
                    const float* g = (const float*)gt;
                    const size_t plane = (size_t)H * W;
                    tgt[j] = make_float3(__ldg(g + pid), __ldg(g + plane + pid), __ldg(g + 2 * plane + pid));
                } else if (GT == 1) {
                    const uint8_t* g = (const uint8_t*)gt + 3 * pid;
                    tgt_u8[j] = (uint32_t)__ldg(g) | ((uint32_t)__ldg(g + 1) << 8) | ((uint32_t)__ldg(g + 2) << 16);
                } else {   // I420: luma plane, then the two quarter-size chroma planes
                    const uint8_t* g = (const uint8_t*)gt;
                    const size_t plane = (size_t)H * W, cid = (size_t)(py_i >> 1) * (W >> 1) + (px_i >> 1);
                    // only loaded here: touching the bytes now would park the warp on DRAM latency
                    // before the batch is even requested
                    raw_y[j] = __ldg(g + pid);
                    raw_u[j] = __ldg(g + plane + cid);
                    raw_v[j] = __ldg(g + plane + (plane >> 2) + cid);
                }
            }
        }
    }

    // ---- everything above is independent of the kernel before this one (the target frame is not
    // its output): under PDL it runs while that kernel drains ----------------------------------
    pdl_wait();
    pdl_release();
    // ranges: the tile's records are [tile_off[t], tile_off[t+1]); slots (tile_off == NULL): the
    // first tile_fill[t] of [t*256, t*256+256) (a fuller tile flagged the step: not reached).
    // The flag and the count are independent loads: one round trip.  (Requesting the first 64
    // records speculatively as well -- their address does not depend on the count -- was measured:
    // no gain, and 25 MB per step of never-written slots pulled from DRAM.)
    const bool slot = tile_off == nullptr;
    // slot mode: a tile's records start at a fixed address, so the first 32 can be requested together
    // with the flag and the count -- one round trip to L2 instead of two.  (Requesting 64 per warp
    // was measured in round 1: no gain, 25 MB per step of never-written slots pulled from DRAM.)
    float4 w0 = make_float4(0.f, 0.f, 0.f, 0.f), w1 = w0, w2 = w0;
#if GS2D_STAGE_BULK
    if (kWarpStage && slot && tid == 0)
        bulk_load_expect(sraw, packed + 3 * ((size_t)t * kSlotStride), 32 * 48, &sbar);
#else
    if (kWarpStage && slot) {
        const float4* rec = packed + 3 * ((size_t)t * kSlotStride + lane);
        w0 = __ldg(rec); w1 = __ldg(rec + 1); w2 = __ldg(rec + 2);
    }
#endif
    const int of = status[kStOverflow];
    const int lo = slot ? t * kSlotStride : tile_off[t];
    const uint32_t fill = slot ? (uint32_t)tile_fill[t] : 0u;          // slots: all pairs << 16 | stored records
    const int cnt = slot ? (int)(fill & 0xffffu) : tile_off[t + 1] - lo;
#if GS2D_STAGE_BULK
    if (kWarpStage && slot) mbar_wait(&sbar, 0);   // every thread passes here once (phase 0), also before an early return
#endif
    if (of != 0) {
        // the workspace overflowed: the step is a no-op.  A render must not hand back stale memory as
        // an image: it is filled with NaN (the operator path reports the overflow lazily)
        if (MODE == 1 && out_img && in_x) {
            const float nanv = __int_as_float(0x7fc00000);
#pragma unroll
            for (int j = 0; j < kPix; ++j) {
                const int py_i = oy + ly0 + 2 * j;
                if (py_i >= H) continue;
                const size_t pid = (size_t)py_i * W + px_i;
                if (out_chw) {
                    const size_t plane = (size_t)H * W;
                    out_img[pid] = nanv; out_img[plane + pid] = nanv; out_img[2 * plane + pid] = nanv;
                } else {
                    out_img[3 * pid] = nanv; out_img[3 * pid + 1] = nanv; out_img[3 * pid + 2] = nanv;
                }
            }
        }
        return;
    }
    // more than 256 and all of them wanted (the reference renders the first 256, Q1): the
    // forward walks the batches in rank order, the backward walks them again; one batch otherwise
    const bool all = ALL && cnt > GS2D_MAX_PER_TILE;
    const int nbat = all ? (cnt + GS2D_MAX_PER_TILE - 1) / GS2D_MAX_PER_TILE : 1;
    GS2D_ASSERT(cnt >= 0 && lo >= 0 && (!slot || cnt <= kSlotStride));
    const bool per_warp = kWarpStage && cnt <= 32;     // uniform over the CTA
    float4* sq = sq_all + (per_warp ? warp * (32 * 3) : 0);
    int nb;
#if GS2D_STAGE_BULK
    if (kWarpStage && slot && per_warp) { w0 = sraw[3 * lane]; w1 = sraw[3 * lane + 1]; w2 = sraw[3 * lane + 2]; }
#endif
    if (per_warp) {
        if (!slot && lane < cnt) {
            const float4* rec = packed + 3 * ((size_t)lo + lane);
            w0 = __ldg(rec); w1 = __ldg(rec + 1); w2 = __ldg(rec + 2);
        }
        stage_tile_warp(w0, w1, w2, cnt, sq, lane);
        nb = cnt;
    } else {
        const int cnt_s = kSqRecords < GS2D_MAX_PER_TILE ? min(cnt, kSqRecords) : cnt;   // (experiment builds truncate)
        nb = stage_tile_sorted<kTileThreads>(packed + 3 * (size_t)lo, cnt_s, sq, sid, tid, all);
    }
    // tile_fill[t] goes back to zero for the next iteration's scatter once every thread of the CTA
    // has read it: after the staging's barriers, or (per-warp staging has none) after the loss barrier
    const bool reset_late = per_warp;
    if (tid == 0 && !reset_late) tile_fill[t] = 0;
    if (tid == 0 && !tile_off && MODE == 0) {
        atomicMax(&status[kStSlotMax], (int)(fill >> 16));   // result unused: a reduction the CTA never waits for
        if (bx == 0 && by == 0)                                    // the update kernel counts the next step into it
            status[kStSlotIsect + ((step + 1) & 1)] = 0;
    }
    if (ids_sorted_dbg && !all)   // parity tests: the order the batch is rendered in
        for (int r = nb + tid; r < cnt; r += kTileThreads) ids_sorted_dbg[lo + r] = -1;

    // ---- forward (forward.cu:590-612) --------------------------------------------------------
    const float px = (float)px_i;
    float py[kPix];
    bool inside[kPix];
#pragma unroll
    for (int j = 0; j < kPix; ++j) {
        py[j] = (float)(oy + ly0 + 2 * j);
        inside[j] = in_x && (oy + ly0 + 2 * j) < H;
    }
    const int wrow_lo = kWarpRows * warp, wrow_hi = kWarpRows * warp + kWarpRows - 1;
    float3 pix[kPix];
    int cur_idx[kPix] = {};
    // Pixels (0,1) and (2,3) of a lane ride in the two lanes of FFMA2 / FMUL2 / FADD2 (sm_100a):
    // the same single-rounding operations per pixel as the scalar form (bit-identical results),
    // half the issue slots for the arithmetic (measured: fit +2.9 %, trained-like +4.2 %, render
    // +3.7..7 %; the same packing of the backward's row pairs cost 31 registers and was 8 % slower).
    // A pixel that fails the test accumulates alpha = 0.  Round 2, late (profiles/r2_ab_fwd_lean.txt): the
    // colours are accumulated by scalar FFMAs again -- the packed accumulators' register pairs were copied at the
    // reconvergence of the row-block branch, six moves per block -- the pass test is two chained predicate sets and
    // one select on the unclamped alpha (gauss_alpha_if_pass), and pixels past the frame's edge are no longer
    // masked per pixel (nothing of theirs is used): 82 -> 74 instructions per record, 46.8 -> 46.0 us at C2,
    // render 23.4 k -> 24.3 k fps.
    f32x2 py2[kPix / 2];
#pragma unroll
    for (int P = 0; P < kPix / 2; ++P) py2[P] = pk2(py[2 * P], py[2 * P + 1]);
    float sr[kPix] = {}, sg_[kPix] = {}, sb[kPix] = {};   // the pixels' accumulated colours
    for (int bat = 0; bat < nbat; ++bat) {
    if (bat) {
        __syncthreads();   // everyone is done with the previous batch
        nb = stage_tile_window<kTileThreads>(packed + 3 * (size_t)lo, cnt, bat * GS2D_MAX_PER_TILE, sq, tid);
    }
    const int lob = lo + bat * GS2D_MAX_PER_TILE;   // sorted index of the batch's first record
    if (ids_sorted_dbg)
        for (int r = tid; r < nb; r += kTileThreads) ids_sorted_dbg[lob + r] = __float_as_int(sq[3 * r + 1].w);
    for (int c = 0; c < nb; c += 32) {
        const int r_lane = c + lane;
        bool touch = false;
        if (r_lane < nb) {
            const uint32_t rect = __float_as_uint(sq[3 * r_lane + 1].z);
            const int y0 = (rect >> 16) & 255, y1 = rect >> 24;
            touch = (y1 >= wrow_lo) && (y0 <= wrow_hi) && (y1 >= y0);   // (an empty rectangle is y0 = 1, y1 = 0)
        }
        unsigned m = __ballot_sync(0xffffffffu, touch);
        while (m) {
            const int k = c + __ffs(m) - 1;
            m &= m - 1;
            const float4 q0 = sq[3 * k];       // x, y, conic.x / 2, conic.y
            const float4 q1 = sq[3 * k + 1];   // conic.z / 2, opacity, rect, id
            const float4 q2 = sq[3 * k + 2];   // r, g, b, tile
            const uint32_t rect = __float_as_uint(q1.z);
            const int y0 = (int)((rect >> 16) & 255) - wrow_lo, y1 = (int)(rect >> 24) - wrow_lo;   // warp-local rows
            const float dx = subf(q0.x, px);
            const float t1 = mulf(q0.z, dx), t3 = mulf(q0.w, dx);   // t1 = (cx * dx) / 2
#pragma unroll
            for (int P = 0; P < kPix / 2; ++P) {
                if (y1 < 4 * P || y0 > 4 * P + 3) continue;   // warp-uniform: rows 4P..4P+3 outside the rectangle
                const f32x2 dy = sub2(bc2(q0.y), py2[P]);
                // reference expression 0.5*(cx*dx*dx + cz*dy*dy) + cy*dx*dy as its build contracts it:
                // fma(dy, cy*dx, 0.5 * fma(dx, cx*dx, (cz*dy)*dy)), the exact factor 0.5 carried by the record
                const f32x2 q = fma2(bc2(dx), bc2(t1), mul2(mul2(bc2(q1.x), dy), dy));
                const f32x2 sg = fma2(dy, bc2(t3), q);
                const f32x2 e = mul2(sg, bc2(-1.4426950216293334961f));
                const f32x2 al = mul2(bc2(q1.y), pk2(ex2_approx(lo2(e)), ex2_approx(hi2(e))));
                // (pixels past the frame's edge are not masked here: nothing of theirs is ever used)
                const float k0 = minf(gauss_alpha_if_pass(lo2(sg), lo2(al)), 1.f);
                const float k1 = minf(gauss_alpha_if_pass(hi2(sg), hi2(al)), 1.f);
                // scalar accumulation (same single roundings): the packed form's register pairs were copied at
                // the reconvergence of the row-block branch, six moves per block
                sr[2 * P] = fmaf_(k0, q2.x, sr[2 * P]); sr[2 * P + 1] = fmaf_(k1, q2.x, sr[2 * P + 1]);
                sg_[2 * P] = fmaf_(k0, q2.y, sg_[2 * P]); sg_[2 * P + 1] = fmaf_(k1, q2.y, sg_[2 * P + 1]);
                sb[2 * P] = fmaf_(k0, q2.z, sb[2 * P]); sb[2 * P + 1] = fmaf_(k1, q2.z, sb[2 * P + 1]);
                if (MODE == 1) {   // a pixel that passed has alpha >= 1/255 (or NaN -> 1 after the min)
                    if (k0 > 0.f) cur_idx[2 * P] = lob + k;
                    if (k1 > 0.f) cur_idx[2 * P + 1] = lob + k;
                }
            }
        }
    }
    }   // batches
#pragma unroll
    for (int P = 0; P < kPix / 2; ++P) {
        pix[2 * P] = make_float3(sr[2 * P], sg_[2 * P], sb[2 * P]);
        pix[2 * P + 1] = make_float3(sr[2 * P + 1], sg_[2 * P + 1], sb[2 * P + 1]);
    }
    if (out_img || final_idx) {
        // operator path (MODE 1, gt = background colour): with no intersection at all the reference
        // returns ones(H, W, 3) * background instead of launching anything (rasterize_sum.py:121-129)
        const float* bg = (MODE == 1 && gt != nullptr && status[kStNumIntersects] < 1) ? (const float*)gt : nullptr;
#pragma unroll
        for (int j = 0; j < kPix; ++j) {
            if (!inside[j]) continue;
            const size_t pid = (size_t)(oy + ly0 + 2 * j) * W + px_i;
            if (out_img) {
                float3 o = pix[j];
                if (MODE == 1 && bg) o = make_float3(bg[0], bg[1], bg[2]);
                if (out_clamp) o = make_float3(fminf(fmaxf(o.x, 0.f), 1.f), fminf(fmaxf(o.y, 0.f), 1.f), fminf(fmaxf(o.z, 0.f), 1.f));
                if (out_chw) {
                    const size_t plane = (size_t)H * W;
                    out_img[pid] = o.x; out_img[plane + pid] = o.y; out_img[2 * plane + pid] = o.z;
                } else {
                    out_img[3 * pid] = o.x; out_img[3 * pid + 1] = o.y; out_img[3 * pid + 2] = o.z;
                }
            }
            if (final_idx) final_idx[pid] = cur_idx[j];
        }
    }
    if (MODE == 1) {
        if (reset_late) {
            __syncthreads();
            if (tid == 0) tile_fill[t] = 0;
        }
        GS2D_TILE_DONE;
    }

    // ---- clamp + L2 loss and its gradient w.r.t. the unclamped sum --------------------------
    float l = 0.f;
#pragma unroll
    for (int j = 0; j < kPix; ++j) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GT == 2) tgt_u8[j] = yuv_to_rgb_packed(raw_y[j], raw_u[j], raw_v[j]);
        if (GT == 1 || GT == 2) tgt[j] = make_float3(u8_over_255(tgt_u8[j] & 255u), u8_over_255((tgt_u8[j] >> 8) & 255u),
                                          u8_over_255((tgt_u8[j] >> 16) & 255u));
        if (inside[j]) {
            const float cr = fminf(fmaxf(pix[j].x, 0.f), 1.f), cg = fminf(fmaxf(pix[j].y, 0.f), 1.f),
                        cb = fminf(fmaxf(pix[j].z, 0.f), 1.f);
            const float dr = cr - tgt[j].x, dg = cg - tgt[j].y, db = cb - tgt[j].z;
            l += dr * dr + dg * dg + db * db;
            // clamp passes the gradient where 0 <= x <= 1 (torch's clamp backward, bounds inclusive),
            // i.e. exactly where clamping left the value unchanged
            v.x = (cr == pix[j].x) ? grad_scale * dr : 0.f;
            v.y = (cg == pix[j].y) ? grad_scale * dg : 0.f;
            v.z = (cb == pix[j].z) ? grad_scale * db : 0.f;
        }
        sv[(ly0 + 2 * j) * GS2D_TILE + lx] = v;   // (the staging's id chunk that aliases sv is dead since its last barrier)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    if (lane == 0) sred[warp] = l;
    __syncthreads();
    if (tid == 0) {
        if (reset_late) tile_fill[t] = 0;
        float s = sred[0];
#pragma unroll
        for (int w = 1; w < kTileThreads / 32; ++w) s += sred[w];
        if (s != 0.f) {
            atomicAdd(loss_acc, (double)s);
            // spatial mode: the loss rides in the pad of gradient row 0 so that one all-reduce of the
            // gradient buffer also sums the ranks' partial losses (pushed reduction: in a word past the rows)
            if (loss_pad) atomicAdd(loss_pad, s);
        }
    }

    // ---- backward (backward.cu:783-859): eight lanes per (tile, Gaussian) pair, four pairs per
    // warp pass.  A lane owns one column of the pair's rectangle (two when it is wider than 8) and
    // walks its rows, so dx and the conic products with dx are per-lane constants; one 7-shuffle
    // transposed reduction leaves the total of value l in lane l of the group, and one reduction
    // instruction per pass adds the four pairs' 8 consecutive floats to their gradient rows.
    const int sub = lane >> 3, l8 = lane & 7;
    for (int bat = nbat - 1; bat >= 0; --bat) {   // the forward left the last batch staged
    if (bat != nbat - 1) {
        __syncthreads();
        nb = stage_tile_window<kTileThreads>(packed + 3 * (size_t)lo, cnt, bat * GS2D_MAX_PER_TILE, sq, tid);
    }
    for (int r0 = warp * 4; r0 < nb; r0 += (kTileThreads / 32) * 4) {
        const int k = min(r0 + sub, nb - 1);
        const float4 q1 = sq[3 * k + 1];
        const uint32_t rect = __float_as_uint(q1.z);
        const int rx0 = rect & 255, rx1 = (rect >> 8) & 255, ry0 = (rect >> 16) & 255, ry1 = rect >> 24;
        const bool live = (r0 + sub < nb) && ry1 >= ry0;
        GS2D_ASSERT(!live || (rx0 <= rx1 && rx1 < GS2D_TILE && ry1 < GS2D_TILE));   // the row walk stays inside sv's 32 rows
        GS2D_ASSERT(!live || (__float_as_int(q1.w) >= 0));
        const int h = live ? ry1 - ry0 + 1 : 0;
        const int ncol = live ? ((rx1 - rx0) >> 3) + 1 : 0;          // column blocks of 8: 1 or 2
        const int iters = h * ncol;
        int max_iters = max(iters, __shfl_xor_sync(0xffffffffu, iters, 8));
        max_iters = max(max_iters, __shfl_xor_sync(0xffffffffu, max_iters, 16));
        if (max_iters == 0) continue;   // no pixel of this tile can pass for any of the four
        const float4 q0 = sq[3 * k], q2 = sq[3 * k + 2];
        const float gx = q0.x, gy = q0.y, cxh = q0.z, cy = q0.w, czh = q1.x, opac = q1.y;   // cxh, czh: conic.x / 2, conic.z / 2
        const float cx = 2.f * cxh, cz = 2.f * czh;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int cb0 = 0; cb0 < 2; ++cb0) {
            int max_h = (cb0 < ncol) ? h : 0;
            max_h = max(max_h, __shfl_xor_sync(0xffffffffu, max_h, 8));
            max_h = max(max_h, __shfl_xor_sync(0xffffffffu, max_h, 16));
            if (max_h == 0) break;
            const int lxp = rx0 + l8 + 8 * cb0;
            const bool col_ok = (cb0 < ncol) && lxp <= rx1;
            const int lxs = min(lxp, GS2D_TILE - 1);                   // a readable column for lanes outside the rectangle
            const float dx = subf(gx, (float)(ox + lxp));
            const float t1 = mulf(cxh, dx), t3 = mulf(cy, dx);
            // Straight-line inner loop: a pixel that fails the reference's test contributes vis = alpha
            // = 0; a lane whose column lies outside its pair's rectangle (col_ok) drops its sums after the walk.  Rows are
            // not masked: the four pairs of a pass walk max_h rows, and a row past a pair's own ry1 is
            // either inside the tile, where the rectangle guarantees that the test fails
            // (gs2d_math.cuh:pair_rect), or past the tile, where v_out reads the zero rows.
            // Per pixel only the moments of v_sigma over dy are accumulated (dx is the lane's
            // constant): S0 = sum v_sigma, S1 = sum v_sigma*dy, S2 = sum v_sigma*dy^2; v_xy and
            // v_conic (backward.cu:828-832) are linear in them and folded in after the column.
            float S0 = 0.f, S1 = 0.f, S2 = 0.f, cr = 0.f, cg = 0.f, cbl = 0.f;
            // dy of the trip's first row; gy - row is exact in fp32 (the result is a multiple of ulp(gy) far
            // below 2^24 ulps), so stepping it by whole rows gives the bits of subf(gy, row) every time
            float dy0 = subf(gy, (float)(oy + ry0));
            const float4* svc = sv + ry0 * GS2D_TILE + lxs;
            // kBwdRows rows per trip: that many independent sigma / ex2 chains in flight per lane
            // (the loop is latency-bound, not issue-bound: profiles/r1 notes)
#pragma unroll kBwdUnroll
            for (int ry = 0; ry < max_h; ry += kBwdRows) {
#pragma unroll
                for (int u = 0; u < kBwdRows; ++u) {
                    const float dy = dy0 - (float)u;
                    const float sg = fmaf_(dy, t3, fmaf_(dx, t1, mulf(mulf(czh, dy), dy)));
                    const float vis = gauss_vis(sg);
                    const float oa = mulf(opac, vis);                                 // alpha before the clamp at 1
                    // gauss_pass(sg, min(oa, 1)) ? oa : 0.  min(oa, 1) < 1/255 <=> oa < 1/255, so the test looks at oa
                    // and ONE select serves the clamped and the unclamped alpha (min(0, 1) = 0); written as two
                    // chained predicate sets + a select (the compiler made it two compare/select pairs).  A lane
                    // outside its pair's columns is not masked per pixel: what it accumulates is dropped after the
                    // column.  28 -> 26 instructions per pixel: 47.9 -> 46.9 us at C2, 121.1 -> 116.0 us with
                    // trained-like Gaussians, same values (profiles/r2_ab_bwd_lean.txt).
                    const float oak = gauss_alpha_if_pass(sg, oa);
                    const float alk = minf(oak, 1.f);
                    const float4 vo = svc[(ry + u) * GS2D_TILE];          // rows past the tile read the zero rows
                    const float v_alpha = q2.x * vo.x + q2.y * vo.y + q2.z * vo.z;   // backward.cu:821-825
                    const float v_sigma = -oak * v_alpha;                             // :827  (-opac * vis * v_alpha)
                    const float t = v_sigma * dy;
                    S0 += v_sigma;
                    S1 += t;
                    S2 = fmaf(t, dy, S2);
                    cr = fmaf(alk, vo.x, cr);                                         // v_rgb    :817
                    cg = fmaf(alk, vo.y, cg);
                    cbl = fmaf(alk, vo.z, cbl);
                }
                dy0 -= (float)kBwdRows;
            }
            if (!col_ok) { S0 = 0.f; S1 = 0.f; S2 = 0.f; cr = 0.f; cg = 0.f; cbl = 0.f; }   // (not a multiply: NaN-proof)
            acc[0] += cx * dx * S0 + cy * S1;                                         // v_xy     :831-832
            acc[1] += cy * dx * S0 + cz * S1;
            acc[2] += 0.5f * dx * dx * S0;                                            // v_conic  :828-830
            acc[3] += 0.5f * dx * S1;
            acc[4] += 0.5f * S2;
            acc[5] += cr; acc[6] += cg; acc[7] += cbl;
        }
        const int gid = __float_as_int(q1.w);
        // pushed reduce-scatter (spatial mode, sharded update): the pair's 32 bytes go straight into the gradient
        // buffer of the rank that OWNS the Gaussian (a reduction without return value: fire and forget over NVLink)
        float* gbase = PUSH ? reinterpret_cast<float*>(__ldg(reinterpret_cast<const unsigned long long*>(push_peers) +
                                                                   (live ? gid / push_per : 0))) : grads;
        float* row = gbase + (size_t)gid * 12;
        if (kRedV4 == 2 || (kRedV4 == 1 && PUSH)) {
            // two 16-byte reductions per pair instead of eight 4-byte ones: a quarter of the packets on NVLink
            float t4[4];
            group8_reduce8_v4(acc, lane, t4);
            if (live && (l8 & 3) == 0 && (t4[0] != 0.f || t4[1] != 0.f || t4[2] != 0.f || t4[3] != 0.f))
                red_add_v4<PUSH>(row + l8, t4);
        } else {
            const float tot = group8_reduce8_t(acc, lane);
            if (live && tot != 0.f) {   // 4 pairs x 8 consecutive floats
                if (PUSH) atomicAdd_system(row + l8, tot);
                else atomicAdd(row + l8, tot);
            }
        }
    }
    }   // batches
#if GS2D_TILE_PERSIST
    }   // tiles
#endif
#undef GS2D_TILE_DONE
}

// ------------------------------------------------------------------------------------------
// K4: projection backward + tanh backward + Adan + gradient re-zeroing
// ------------------------------------------------------------------------------------------
// the ranks' partial losses: one word per peer, all requests in flight together (a plain `acc += load` loop
// waits for each NVLink round trip in turn -- on the ONE thread that reports the step, which the whole kernel
// then waits for: 2 us per peer)
__device__ __forceinline__ float sum_peer_words(const float* const* __restrict__ peers, int num_peers, size_t at) {
    float acc = 0.f;
    for (int r0 = 0; r0 < num_peers; r0 += 8) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = (r0 + k < num_peers) ? __ldcg(peers[r0 + k] + at) : 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) acc += v[k];
    }
    return acc;
}

// one 16-byte row segment summed over all ranks by the NVLink switch (NVLS)
__device__ __forceinline__ float4 multimem_sum4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ float multimem_sum1(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}

struct AdanHyper {
    float b1, b2, b3;
    float step_size;        // lr / (1 - b1^t)
    float step_size_diff;   // lr * b2 / (1 - b2^t)
    float inv_bc3_sqrt;     // 1 / sqrt(1 - b3^t)
    float eps;
    float decay_div;        // 1 + lr * weight_decay
    int first;              // t == 1: previous gradient := gradient
    float offdiag;          // 2: the reference's projection backward; 1: analytic gradient
};

__device__ __forceinline__ float adan_one(float p, float g, float& m, float& nn, float& d, float& prev,
                                          const AdanHyper& h) {
    const float diff = h.first ? 0.f : g - prev;
    m = m * h.b1 + (1.f - h.b1) * g;
    d = d * h.b2 + (1.f - h.b2) * diff;
    const float u = g + h.b2 * diff;
    nn = nn * h.b3 + (1.f - h.b3) * u * u;
    const float denom = sqrtf(nn) * h.inv_bc3_sqrt + h.eps;
    prev = g;
    float q = p - h.step_size * (m / denom) - h.step_size_diff * (d / denom);
    // x / 1 == x: no weight decay (the reference's setting), no division.  The empty asm keeps the compiler from
    // turning the (uniform) branch into a select that computes the IEEE division anyway -- it did: 4 % of the
    // update kernel's instructions (profiles/r2_fit_update_hotspots.txt)
    if (h.decay_div != 1.f) {
        asm volatile("" ::: "memory");
        q = q / h.decay_div;
    }
    return q;
}

// PACKED: the parameters live in one fp32 [n][8] buffer (xyz 2 | chol 3 | feat 3; gs2d_fit_options.params8),
// passed as xyz_raw -- two aligned 16-byte loads and stores per Gaussian where the reference's AoS tensors
// (bindings.cu:456-465 casts them to float2 / float3) take 8 scalar ones; chol_raw / feat are then unused
// and the sharded update's route holds the peers' packed buffers in rt.xyz.
template <bool PACKED>
__global__ void __launch_bounds__(kFitThreads)
k_fit_update(int n, float* __restrict__ xyz_raw, float* __restrict__ chol_raw, float* __restrict__ feat,
             float* __restrict__ adan, float* __restrict__ grads, float half_w, float half_h, int tx, int ty,
             AdanHyper h, int32_t* status, const double* loss_this, double* loss_next,
             float* __restrict__ grads_out, int cap, int next_step, int32_t* tile_count,
             int32_t* __restrict__ tile_off, gs2d_fit_report* report_host, int row0, int row1, int loss_in_grads,
             const float* const* __restrict__ peer_grads, int num_peers, float* __restrict__ grads_zero,
             float* __restrict__ rgb_w, float* __restrict__ adan_w, float* __restrict__ rgb_w_grad_out,
             int skip_update, const float* __restrict__ mc_grads, int slots, float4* __restrict__ packed, int W,
             int H, int shard_begin, int shard_end, Route rt) {
    __shared__ int sh[8];
    __shared__ int s_last;
    __shared__ float s_par[PACKED ? 1 : 8 * kFitThreads];
    float4* p8 = reinterpret_cast<float4*>(xyz_raw);
    // this launch updates Gaussians [shard_begin, shard_end) -- [0, n) unless the update is sharded
    // over the ranks; all arrays are indexed by the global Gaussian index
    const int gtid = blockIdx.x * kFitThreads + threadIdx.x;
    const int i = shard_begin + gtid;
    // ---- prologue: parameters and optimizer state were last written by the PREVIOUS update kernel,
    // not by the tile kernel this one follows -- under PDL all of this (the bulk of the kernel's
    // memory traffic and both tanh) runs while the tile kernel's last wave drains ----------------
    Proj2d p; float2 m; float3 L;
    float par[8], am[8], an[8], ad[8], ap[8];
    float cw = 1.f, wm = 0.f, wn = 0.f, wd = 0.f, wp = 0.f;
    const size_t plane = (size_t)n * 8;
    float4* sm = reinterpret_cast<float4*>(adan + (size_t)i * 8);
    float4* sn = reinterpret_cast<float4*>(adan + plane + (size_t)i * 8);
    float4* sd = reinterpret_cast<float4*>(adan + 2 * plane + (size_t)i * 8);
    float4* sp = reinterpret_cast<float4*>(adan + 3 * plane + (size_t)i * 8);
    if (i < shard_end) {
        if (PACKED) {
            *reinterpret_cast<float4*>(par) = p8[2 * (size_t)i];
            *reinterpret_cast<float4*>(par + 4) = p8[2 * (size_t)i + 1];
            m = make_float2(tanhf(par[0]), tanhf(par[1]));
            L = make_float3(addf(par[2], 0.5f), addf(par[3], 0.f), addf(par[4], 0.5f));
            project_chol(m.x, m.y, L.x, L.y, L.z, half_w, half_h, tx, ty, p);
        } else {
            fit_project(i, xyz_raw, chol_raw, 1, half_w, half_h, tx, ty, p, m, L, QuantConst());
            par[0] = xyz_raw[2 * i]; par[1] = xyz_raw[2 * i + 1];
            par[2] = chol_raw[3 * i]; par[3] = chol_raw[3 * i + 1]; par[4] = chol_raw[3 * i + 2];
            par[5] = feat[3 * i]; par[6] = feat[3 * i + 1]; par[7] = feat[3 * i + 2];
        }
        if (!skip_update) {
            *reinterpret_cast<float4*>(am) = sm[0]; *reinterpret_cast<float4*>(am + 4) = sm[1];
            *reinterpret_cast<float4*>(an) = sn[0]; *reinterpret_cast<float4*>(an + 4) = sn[1];
            *reinterpret_cast<float4*>(ad) = sd[0]; *reinterpret_cast<float4*>(ad + 4) = sd[1];
            *reinterpret_cast<float4*>(ap) = sp[0]; *reinterpret_cast<float4*>(ap + 4) = sp[1];
        }
        if (rgb_w) {
            cw = rgb_w[i];
            if (!skip_update) { wm = adan_w[i]; wn = adan_w[(size_t)n + i]; wd = adan_w[2 * (size_t)n + i]; wp = adan_w[3 * (size_t)n + i]; }
        }
    }
    pdl_wait();
    pdl_release();
    // a flag for a LATER step (slot mode: another block may be raising it right now for the step
    // being scattered below) must not stop this step's update half way
    const int of = status[kStOverflow];
    const bool dead = of != 0 && abs(of) < next_step;
    if (gtid == 0 && report_host) {
        // the step's result straight into pinned host memory (no copy to issue, nothing to wait
        // for but the kernel itself); written before the scan below replaces status[0]
        double ls = *loss_this;
        if (loss_in_grads) {
            if (mc_grads) {
                ls = (double)multimem_sum1(mc_grads + 9);
            } else if (num_peers > 0) {
                ls = (double)sum_peer_words(peer_grads, num_peers, 9);
            } else {
                ls = (double)grads[9];
            }
        }
        report_host->loss_sum = ls;
        report_host->num_intersects = slots ? status[kStSlotIsect + ((next_step - 1) & 1)] : status[kStNumIntersects];
        report_host->overflow_step = of;
        report_host->max_per_tile = slots ? status[kStSlotMax] : status[kStMaxPerTile];
        if (slots) status[kStSlotMax] = 0;   // the tile kernel of the next step raises it again
        __threadfence_system();
        report_host->step = next_step - 1;   // written last: the host treats it as the valid flag
        __threadfence_system();
    }
    if (dead) return;
    if (gtid == 0 && loss_next) *loss_next = 0.0;
    if (rt.R > 0 && grads_zero) {
        // sharded: this rank reads (and would re-zero) only its shard's rows, but its tile kernel
        // adds into every row of its own buffer -- clear the whole spare buffer for the step after next
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        float4* zb = reinterpret_cast<float4*>(grads_zero);
        for (size_t q = gtid; q < (size_t)n * 3; q += (size_t)gridDim.x * kFitThreads) zb[q] = z;
    }
    int claimed = 0;
    GS2D_ASSERT(shard_begin >= 0 && shard_end <= n);
    if (i < shard_end) {
        float4 g0, g1;   // v_xy(2) v_conic(3) v_rgb(3) | v_opacity(1) pad(3) unused here
        if (mc_grads) {
            // spatial mode over the switch: one load returns the row summed over all ranks
            g0 = multimem_sum4(mc_grads + (size_t)i * 12);
            g1 = multimem_sum4(mc_grads + (size_t)i * 12 + 4);
        } else if (num_peers > 0) {
            // spatial mode over peer memory: the all-reduce IS this read -- every rank sums the
            // ranks' gradient rows (NVLink peer loads) in rank order, so all ranks compute
            // bit-identical updates with no collective kernel and no extra pass over the rows
            g0 = make_float4(0.f, 0.f, 0.f, 0.f);
            g1 = g0;
            // up to eight peers' rows in flight at once: one NVLink round trip, not one per rank
            for (int r0 = 0; r0 < num_peers; r0 += 8) {
                float4 a[8], b[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (r0 + k < num_peers) {
                        const float4* prow = reinterpret_cast<const float4*>(peer_grads[r0 + k] + (size_t)i * 12);
                        a[k] = __ldcg(prow); b[k] = __ldcg(prow + 1);
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (r0 + k < num_peers) {
                        g0.x += a[k].x; g0.y += a[k].y; g0.z += a[k].z; g0.w += a[k].w;
                        g1.x += b[k].x; g1.y += b[k].y; g1.z += b[k].z; g1.w += b[k].w;
                    }
                }
            }
        } else {
            const float4* grow = reinterpret_cast<const float4*>(grads + (size_t)i * 12);
            g0 = grow[0]; g1 = grow[1];
        }
        if (rt.R == 0) {   // re-zero a gradient row for a later step: this step's buffer, or (peer exchange) the
            // OTHER buffer of the pair, which no peer can still be reading (see SpatialFrameFit)
            float4* zrow = reinterpret_cast<float4*>((grads_zero ? grads_zero : grads) + (size_t)i * 12);
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            zrow[0] = z; zrow[1] = z; zrow[2] = z;
        }
        float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, g1.y, g1.z, g1.w};   // xyz(2) chol(3) feat(3)
        float gw = 0.f;
        if (rgb_w) {   // colours = feat * w: d feat = d colour * w, d w = <d colour, feat>
            gw = g1.y * par[5] + g1.z * par[6] + g1.w * par[7];
            g[5] = g1.y * cw; g[6] = g1.z * cw; g[7] = g1.w * cw;
            if (rgb_w_grad_out) rgb_w_grad_out[i] = gw;
        }
        if (p.ok && (int)p.radius > 0) {   // backward2d.cu:23-25
            const float3 vc = conic_vjp(p.conic, make_float3(g0.z, g0.w, g1.x));
            const float3 vl = chol_vjp(L.x, L.y, L.z, vc, h.offdiag);
            const float vmx = g0.x * half_w, vmy = g0.y * half_h;            // backward2d.cu:48-49
            g[0] = vmx * (1.f - m.x * m.x);                                   // d tanh
            g[1] = vmy * (1.f - m.y * m.y);
            g[2] = vl.x; g[3] = vl.y; g[4] = vl.z;
        }
        if (grads_out) {   // parity tests: the parameter gradients autograd would hand to the optimizer
#pragma unroll
            for (int k = 0; k < 8; ++k) grads_out[(size_t)i * 8 + k] = g[k];
        }
        if (skip_update) return;   // uniform over the grid: nobody reaches the scan below
#pragma unroll
        for (int k = 0; k < 8; ++k) par[k] = adan_one(par[k], g[k], am[k], an[k], ad[k], ap[k], h);
        sm[0] = *reinterpret_cast<float4*>(am); sm[1] = *reinterpret_cast<float4*>(am + 4);
        sn[0] = *reinterpret_cast<float4*>(an); sn[1] = *reinterpret_cast<float4*>(an + 4);
        sd[0] = *reinterpret_cast<float4*>(ad); sd[1] = *reinterpret_cast<float4*>(ad + 4);
        sp[0] = *reinterpret_cast<float4*>(ap); sp[1] = *reinterpret_cast<float4*>(ap + 4);
        if (PACKED) {
            const float4 a = make_float4(par[0], par[1], par[2], par[3]), b = make_float4(par[4], par[5], par[6], par[7]);
            p8[2 * (size_t)i] = a;
            p8[2 * (size_t)i + 1] = b;
            // sharded update: the all-gather is these stores into the peers' buffers (NVLink, fire and
            // forget; read after the barrier that follows this kernel) -- 32 contiguous bytes per
            // Gaussian, so a warp fills 1 KB of consecutive lines in each peer
            for (int q = 0; q < rt.R; ++q) {
                float4* pq = reinterpret_cast<float4*>(rt.xyz[q]);
                if (pq == p8) continue;
                pq[2 * (size_t)i] = a;
                pq[2 * (size_t)i + 1] = b;
            }
        } else {
            xyz_raw[2 * i] = par[0]; xyz_raw[2 * i + 1] = par[1];
            chol_raw[3 * i] = par[2]; chol_raw[3 * i + 1] = par[3]; chol_raw[3 * i + 2] = par[4];
            feat[3 * i] = par[5]; feat[3 * i + 1] = par[6]; feat[3 * i + 2] = par[7];
        }
        if (!PACKED && rt.R > 0) {   // sharded update: staged for the block's stores into the peers' tensors (below)
            s_par[2 * threadIdx.x] = par[0]; s_par[2 * threadIdx.x + 1] = par[1];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                s_par[2 * kFitThreads + 3 * threadIdx.x + k] = par[2 + k];
                s_par[5 * kFitThreads + 3 * threadIdx.x + k] = par[5 + k];
            }
        }
        if (rgb_w) {
            cw = adan_one(cw, gw, wm, wn, wd, wp, h);
            rgb_w[i] = cw;
            adan_w[i] = wm; adan_w[(size_t)n + i] = wn; adan_w[2 * (size_t)n + i] = wd; adan_w[3 * (size_t)n + i] = wp;
        }
        // the next iteration's binning starts here: project the updated parameters and count tiles
        // (same arithmetic as k_fit_count / k_fit_scatter, so the ranges match what scatter fills);
        // not when the update is sharded: each rank bins its band from the gathered parameters
        if (rt.R > 0 && rt.work) {
            // sharded update with remote scatter: the records of this Gaussian go straight to the ranks
            // whose bands it touches
            Proj2d pn;
            project_chol(tanhf(par[0]), tanhf(par[1]), addf(par[2], 0.5f), addf(par[3], 0.f), addf(par[4], 0.5f),
                         half_w, half_h, tx, ty, pn);
            claimed = scatter_records_remote(pn, i, 1.f, mulf(par[5], cw), mulf(par[6], cw), mulf(par[7], cw), tx, W, H,
                                             status, next_step, rt);
        }
        if (rt.R == 0) {
            Proj2d pn;
            project_chol(tanhf(par[0]), tanhf(par[1]), addf(par[2], 0.5f), addf(par[3], 0.f), addf(par[4], 0.5f),
                         half_w, half_h, tx, ty, pn);
            if (slots) {
                // ... and with fixed slots per tile there is nothing to count or scan first: the records
                // of the next iteration are written right here (tile_count is the tiles' fill level)
                claimed = scatter_records<true>(pn, i, 1.f, mulf(par[5], cw), mulf(par[6], cw), mulf(par[7], cw), tx,
                                                row0, row1, W, H, nullptr, tile_count, packed, status, next_step, Route(),
                                                (slots & 2) != 0);
            } else {
                count_tiles(pn, tx, row0, row1, tile_count);
            }
        }
    }
    if (skip_update) return;
    if (rt.R > 0 && rt.work) {
        block_add(claimed, &status[kStSlotIsect + (next_step & 1)], sh);   // (this shard's pairs, over all bands)
        if (PACKED) return;
    }
    if (PACKED && rt.R > 0) return;
    if (rt.R > 0) {
        // sharded update: the all-gather is these stores (NVLink, fire and forget; the peers read
        // them after the barrier that follows this kernel).  A block's Gaussians are contiguous in
        // each of the three tensors, so every store instruction of a warp fills one 128-byte line.
        __syncthreads();
        const size_t i0 = (size_t)shard_begin + (size_t)blockIdx.x * kFitThreads;
        const int c = min(kFitThreads, shard_end - (int)i0);
        for (int q = 0; q < rt.R; ++q) {
            float* px = rt.xyz[q];
            if (px == xyz_raw) continue;   // this rank's own copy was written above
            float* pc = rt.chol[q]; float* pf = rt.feat[q];
            for (int k = threadIdx.x; k < 2 * c; k += kFitThreads) px[2 * i0 + k] = s_par[k];
            for (int k = threadIdx.x; k < 3 * c; k += kFitThreads) {
                pc[3 * i0 + k] = s_par[2 * kFitThreads + k];
                pf[3 * i0 + k] = s_par[5 * kFitThreads + k];
            }
        }
        return;
    }
    if (slots) {
        block_add(claimed, &status[kStSlotIsect + (next_step & 1)], sh);
        return;
    }
    scan_tiles_by_last_block(tx, ty, cap, next_step, tile_count, tile_off, status, sh, &s_last);
}

// The same update with FOUR lanes per Gaussian (packed [n][8] rows only).  One thread per Gaussian is one
// dependent chain of ~1 500 instructions on ~10 warps per SM: latency-bound (round 1: 22 % issue-active).
// Here lane q of a quad runs Adan on parameters 2q, 2q+1 (their state rows are read and written as float2:
// four lanes cover a Gaussian's 32 bytes), every lane recomputes the cheap shared parts (projection of the
// old parameters for the backward, of the new ones for the binning), the new parameters are exchanged by
// shuffles inside the quad, and the next step's tiles are shared out over the four lanes (tile j goes to
// lane j mod 4) so that the slot claims of a Gaussian are in flight together.  Same arithmetic per element,
// hence the same bits as k_fit_update<true>; the learnable colour weight rides on lane 0.
__device__ __forceinline__ float pick4(int q, float a, float b, float c, float d) {
    return q == 0 ? a : (q == 1 ? b : (q == 2 ? c : d));
}

#ifndef GS2D_UPDATE_LANES
#define GS2D_UPDATE_LANES 2   // measured at C2 (adan stage): 1 lane 18.4 us, 4 lanes 16.1, 2 lanes 14.8 (one wave at 85 registers)
#endif
constexpr int kUpdateLanes = GS2D_UPDATE_LANES;
#ifndef GS2D_QUAD_MIN_BLOCKS
#define GS2D_QUAD_MIN_BLOCKS 6   // 85 registers: 24 warps per SM (measured: 16.6 us against 18.8 at the compiler's own 96)
#endif
// MULTI = false: the single-GPU instantiation carries none of the cross-rank paths (peer / switch reads of the
// gradient rows, parameter stores into peers, remote scatter).  The kernel executes every instruction once
// per warp, so its time is instruction FETCH as much as anything (round 2: 44 % of the one-thread kernel's
// stall samples were "no instruction", 7 300 SASS instructions): less code is less time.
// LANES = 4 or 2 lanes per Gaussian: lane q owns parameters [q*K, q*K + K), K = 8 / LANES
template <bool MULTI, int LANES>
__global__ void __launch_bounds__(kFitThreads, GS2D_QUAD_MIN_BLOCKS)
k_fit_update_quad(int n, float* __restrict__ params8, float* __restrict__ adan, float* __restrict__ grads, float half_w,
                  float half_h, int tx, int ty, AdanHyper h, int32_t* status, const double* loss_this, double* loss_next,
                  float* __restrict__ grads_out, int cap, int next_step, int32_t* tile_count,
                  int32_t* __restrict__ tile_off, gs2d_fit_report* report_host, int row0, int row1, int loss_in_grads,
                  const float* const* __restrict__ peer_grads, int num_peers, float* __restrict__ grads_zero,
                  float* __restrict__ rgb_w, float* __restrict__ adan_w, float* __restrict__ rgb_w_grad_out,
                  int skip_update, const float* __restrict__ mc_grads, int slots, float4* __restrict__ packed, int W,
                  int H, int shard_begin, int shard_end, Route rt) {
    __shared__ int sh[8];
    __shared__ int s_last;
    const int gtid = blockIdx.x * kFitThreads + threadIdx.x;
    constexpr int K = 8 / LANES;                  // parameters per lane
    static_assert(LANES == 2 || LANES == 4, "two or four lanes per Gaussian");
    const int q = gtid & (LANES - 1);             // lane of the group: parameters q*K .. q*K + K - 1
    const int i = shard_begin + gtid / LANES;
    const bool act = i < shard_end;
    const unsigned full = 0xffffffffu;
    const float4* p8 = reinterpret_cast<const float4*>(params8);
    // ---- prologue (runs under the tile kernel's tail, see k_fit_update) ----------------------------
    Proj2d p; float2 m = make_float2(0.f, 0.f); float3 L = make_float3(1.f, 0.f, 1.f);
    float par[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    __align__(16) float am[K] = {}, an[K] = {}, ad[K] = {}, ap[K] = {};
    typedef typename std::conditional<K == 2, float2, float4>::type VecK;   // a lane's K floats as one access
    float cw = 1.f, wm = 0.f, wn = 0.f, wd = 0.f, wp = 0.f;
    const size_t plane = (size_t)n * 8;
    const size_t so = (size_t)(act ? i : 0) * 8 + K * q;   // this lane's share of a state / parameter row
    p.ok = false; p.radius = 0.f; p.nth = 0; p.conic = make_float3(0.f, 0.f, 0.f);
    if (act) {
        *reinterpret_cast<float4*>(par) = p8[2 * (size_t)i];
        *reinterpret_cast<float4*>(par + 4) = p8[2 * (size_t)i + 1];
        m = make_float2(tanhf(par[0]), tanhf(par[1]));
        L = make_float3(addf(par[2], 0.5f), addf(par[3], 0.f), addf(par[4], 0.5f));
        project_chol(m.x, m.y, L.x, L.y, L.z, half_w, half_h, tx, ty, p);
        if (!skip_update) {
            *reinterpret_cast<VecK*>(am) = *reinterpret_cast<const VecK*>(adan + so);
            *reinterpret_cast<VecK*>(an) = *reinterpret_cast<const VecK*>(adan + plane + so);
            *reinterpret_cast<VecK*>(ad) = *reinterpret_cast<const VecK*>(adan + 2 * plane + so);
            *reinterpret_cast<VecK*>(ap) = *reinterpret_cast<const VecK*>(adan + 3 * plane + so);
        }
        if (rgb_w) {
            cw = rgb_w[i];
            if (!skip_update && q == 0) { wm = adan_w[i]; wn = adan_w[(size_t)n + i]; wd = adan_w[2 * (size_t)n + i]; wp = adan_w[3 * (size_t)n + i]; }
        }
    }
    pdl_wait();
    pdl_release();
    // the local gradient row is requested together with the status word (one round trip, not two)
    float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;   // v_xy(2) v_conic(3) v_rgb(3) | v_opacity(1) pad(3)
    const bool push = MULTI && (loss_in_grads & 2);          // the peers' tile kernels added into THIS rank's rows
    const bool local_row = !MULTI || push || (!mc_grads && num_peers == 0);
    if (act && local_row) {
        const float4* grow = reinterpret_cast<const float4*>(grads + (size_t)i * 12);
        g0 = grow[0]; g1 = grow[1];
    }
    const int of = status[kStOverflow];
    const bool dead = of != 0 && abs(of) < next_step;
    if (gtid == 0 && report_host) {
        double ls = *loss_this;
        if (loss_in_grads & 1) {
            if (MULTI && mc_grads) {
                ls = (double)multimem_sum1(mc_grads + 9);
            } else if (MULTI && num_peers > 0) {
                // pushed reduction: the ranks' partial losses sit past the rows, in one of two words by step parity
                const size_t pad = push ? (size_t)n * 12 + 16 * ((next_step - 1) & 1) : 9;
                ls = (double)sum_peer_words(peer_grads, num_peers, pad);
            } else {
                ls = (double)grads[9];
            }
        }
        report_host->loss_sum = ls;
        report_host->num_intersects = slots ? status[kStSlotIsect + ((next_step - 1) & 1)] : status[kStNumIntersects];
        report_host->overflow_step = of;
        report_host->max_per_tile = slots ? status[kStSlotMax] : status[kStMaxPerTile];
        if (slots) status[kStSlotMax] = 0;
        __threadfence_system();
        report_host->step = next_step - 1;
        __threadfence_system();
    }
    if (dead) return;                              // uniform over the grid (see k_fit_update)
    if (gtid == 0 && loss_next) *loss_next = 0.0;
    // (the next step's loss word: its last readers were the update kernels of the step before this one)
    if (push && gtid == 0) grads[(size_t)n * 12 + 16 * (next_step & 1)] = 0.f;
    if (MULTI && rt.R > 0 && grads_zero && !push) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        float4* zb = reinterpret_cast<float4*>(grads_zero);
        for (size_t k = gtid; k < (size_t)n * 3; k += (size_t)gridDim.x * kFitThreads) zb[k] = z;
    }
    int claimed = 0;
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float gw = 0.f;
    if (act) {
        if (MULTI && mc_grads) {   // (the group's lanes read the same row)
            g0 = multimem_sum4(mc_grads + (size_t)i * 12);
            g1 = multimem_sum4(mc_grads + (size_t)i * 12 + 4);
        } else if (MULTI && num_peers > 0 && !push) {
            g0 = make_float4(0.f, 0.f, 0.f, 0.f);
            g1 = g0;
            for (int r0 = 0; r0 < num_peers; r0 += 4) {
                float4 a[4], b[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (r0 + k < num_peers) {
                        const float4* prow = reinterpret_cast<const float4*>(peer_grads[r0 + k] + (size_t)i * 12);
                        a[k] = __ldcg(prow); b[k] = __ldcg(prow + 1);
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (r0 + k < num_peers) {
                        g0.x += a[k].x; g0.y += a[k].y; g0.z += a[k].z; g0.w += a[k].w;
                        g1.x += b[k].x; g1.y += b[k].y; g1.z += b[k].z; g1.w += b[k].w;
                    }
                }
            }
        }
        g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
        if (rgb_w) {
            gw = g1.y * par[5] + g1.z * par[6] + g1.w * par[7];
            g[5] = g1.y * cw; g[6] = g1.z * cw; g[7] = g1.w * cw;
            if (rgb_w_grad_out && q == 0) rgb_w_grad_out[i] = gw;
        }
        if (p.ok && (int)p.radius > 0) {   // backward2d.cu:23-25
            const float3 vc = conic_vjp(p.conic, make_float3(g0.z, g0.w, g1.x));
            const float3 vl = chol_vjp(L.x, L.y, L.z, vc, h.offdiag);
            const float vmx = g0.x * half_w, vmy = g0.y * half_h;            // backward2d.cu:48-49
            g[0] = vmx * (1.f - m.x * m.x);                                   // d tanh
            g[1] = vmy * (1.f - m.y * m.y);
            g[2] = vl.x; g[3] = vl.y; g[4] = vl.z;
        }
    }
    // every lane of the warp has loaded its row: now the row may be re-zeroed (16 bytes per lane).
    // The barrier orders the group's loads before the stores, whichever lane is ahead.
    __syncwarp();
    if (act && (!MULTI || rt.R == 0 || push)) {
        float4* zrow = reinterpret_cast<float4*>(((grads_zero && !push) ? grads_zero : grads) + (size_t)i * 12);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        if (LANES == 4) { if (q < 3) zrow[q] = z; }
        else { zrow[q] = z; if (q == 0) zrow[2] = z; }
    }
    __align__(16) float gm[K], pm[K];              // this lane's gradients and parameters
#pragma unroll
    for (int k = 0; k < K; ++k) {
        gm[k] = LANES == 4 ? pick4(q, g[k], g[K + k], g[2 * K + k], g[3 * K + k]) : (q == 0 ? g[k] : g[K + k]);
        pm[k] = LANES == 4 ? pick4(q, par[k], par[K + k], par[2 * K + k], par[3 * K + k]) : (q == 0 ? par[k] : par[K + k]);
    }
    if (act && grads_out) *reinterpret_cast<VecK*>(grads_out + (size_t)i * 8 + K * q) = *reinterpret_cast<VecK*>(gm);
    if (skip_update) return;                       // uniform over the grid
    if (act) {
#pragma unroll
        for (int k = 0; k < K; ++k) pm[k] = adan_one(pm[k], gm[k], am[k], an[k], ad[k], ap[k], h);
        *reinterpret_cast<VecK*>(adan + so) = *reinterpret_cast<VecK*>(am);
        *reinterpret_cast<VecK*>(adan + plane + so) = *reinterpret_cast<VecK*>(an);
        *reinterpret_cast<VecK*>(adan + 2 * plane + so) = *reinterpret_cast<VecK*>(ad);
        *reinterpret_cast<VecK*>(adan + 3 * plane + so) = *reinterpret_cast<VecK*>(ap);
        *reinterpret_cast<VecK*>(params8 + so) = *reinterpret_cast<VecK*>(pm);
        for (int z = 0; MULTI && z < rt.R; ++z) {  // sharded update: the all-gather is these NVLink stores
            float* pq = rt.xyz[z];
            if (pq != params8) *reinterpret_cast<VecK*>(pq + so) = *reinterpret_cast<VecK*>(pm);
        }
        if (rgb_w && q == 0) {
            cw = adan_one(cw, gw, wm, wn, wd, wp, h);
            rgb_w[i] = cw;
            adan_w[i] = wm; adan_w[(size_t)n + i] = wn; adan_w[2 * (size_t)n + i] = wd; adan_w[3 * (size_t)n + i] = wp;
        }
    }
    // the quad's eight new parameters (and the new colour weight) to every lane of it
    float np[8];
#pragma unroll
    for (int l = 0; l < LANES; ++l)
#pragma unroll
        for (int k = 0; k < K; ++k) np[l * K + k] = __shfl_sync(full, pm[k], l, LANES);
    cw = __shfl_sync(full, cw, 0, LANES);
    if (act && (!MULTI || rt.R == 0 || rt.work)) {
        Proj2d pn;
        project_chol(tanhf(np[0]), tanhf(np[1]), addf(np[2], 0.5f), addf(np[3], 0.f), addf(np[4], 0.5f), half_w, half_h,
                     tx, ty, pn);
        const float cr = mulf(np[5], cw), cg = mulf(np[6], cw), cb = mulf(np[7], cw);
        if (MULTI && rt.R > 0) {                   // remote scatter (option): one lane walks the owners in order
            if (q == 0) claimed = scatter_records_remote(pn, i, 1.f, cr, cg, cb, tx, W, H, status, next_step, rt);
        } else if (slots) {
            claimed = scatter_records<true, 8 / LANES>(pn, i, 1.f, cr, cg, cb, tx, row0, row1, W, H, nullptr, tile_count,
                                                       packed, status, next_step, Route(), (slots & 2) != 0, q, LANES);
        } else {
            count_tiles(pn, tx, row0, row1, tile_count, q, LANES);
        }
    }
    if (MULTI && rt.R > 0) {
        if (rt.work) block_add(claimed, &status[kStSlotIsect + (next_step & 1)], sh);
        return;
    }
    if (slots) {
        block_add(claimed, &status[kStSlotIsect + (next_step & 1)], sh);
        return;
    }
    scan_tiles_by_last_block(tx, ty, cap, next_step, tile_count, tile_off, status, sh, &s_last);
}

// Cross-rank barrier on the stream (include/gs2d.h, gs2d_grad_exchange.barrier_pads): one warp, lane
// q signals rank q and waits for rank q.  The kernels before it on this stream have completed, so
// their writes (local REDs, stores into peer memory) precede the fence; the release / acquire pair
// at system scope hands them to whoever passes the barrier on the other side.  The kernel that
// follows may be launched programmatically: its prologue runs under the wait.
__global__ void __launch_bounds__(32)
k_rank_barrier(uint32_t* const* __restrict__ pads, int R, int rank, uint32_t epoch) {
    // launched programmatically itself: the warp is already resident when the kernel before it drains
    // (its launch latency is off the step's critical path); nothing is signalled before that kernel's
    // writes are complete and visible
    pdl_wait();
    pdl_release();
    const int q = threadIdx.x;
    if (q >= R) return;
    __threadfence_system();
    uint32_t* dst = pads[q] + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
    const uint32_t* src = pads[rank] + q;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    uint32_t v;
    for (;;) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
        if ((int32_t)(v - epoch) >= 0) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 10000000000ull) __trap();   // 10 s: a peer is gone
    }
}

// wire format -> the fp32 parameter tensors of the frame model (what forward_quantize feeds the two
// operators BEFORE tanh / bound: fp16-rounded _xyz, code*scale+beta, e0+e1): for a decoder that goes
// on fitting (P-frames start from the decoded previous frame) and for the parity tests
__global__ void __launch_bounds__(kFitThreads)
k_dequantize(int n, const uint16_t* __restrict__ xyz_half, const uint8_t* __restrict__ chol_code,
             const uint8_t* __restrict__ color_index, QuantConst qc, float* __restrict__ xyz_raw,
             float* __restrict__ chol_raw, float* __restrict__ features) {
    const int i = blockIdx.x * kFitThreads + threadIdx.x;
    if (i >= n) return;
    xyz_raw[2 * i] = __half2float(__ushort_as_half(xyz_half[2 * i]));
    xyz_raw[2 * i + 1] = __half2float(__ushort_as_half(xyz_half[2 * i + 1]));
#pragma unroll
    for (int c = 0; c < 3; ++c)
        chol_raw[3 * i + c] = addf(mulf((float)chol_code[3 * i + c], qc.scale[c]), qc.beta[c]);
    const float3 col = fit_colour(i, reinterpret_cast<const float*>(color_index), nullptr, 2, qc);
    features[3 * i] = col.x; features[3 * i + 1] = col.y; features[3 * i + 2] = col.z;
}

// the loader as a pass of its own (drop-in path: the model files want an RGB tensor): one thread
// per 2x2 block -- two 2-byte luma loads, one U, one V, two 6-byte row segments out
__global__ void __launch_bounds__(kFitThreads)
k_i420_to_rgb(int H, int W, const uint8_t* __restrict__ yuv, uint8_t* __restrict__ rgb) {
    const int bw = W >> 1, bh = H >> 1;
    const int b = blockIdx.x * kFitThreads + threadIdx.x;
    if (b >= bw * bh) return;
    const int by = b / bw, bx = b - by * bw;
    const size_t plane = (size_t)H * W;
    const int u = __ldg(yuv + plane + b), v = __ldg(yuv + plane + (plane >> 2) + b);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const size_t pid = (size_t)(2 * by + r) * W + 2 * bx;
        const uint16_t yy = __ldg(reinterpret_cast<const uint16_t*>(yuv + pid));   // pid is even
        const uint32_t a = yuv_to_rgb_packed(yy & 255, u, v), c = yuv_to_rgb_packed(yy >> 8, u, v);
        uint16_t* o = reinterpret_cast<uint16_t*>(rgb + 3 * pid);                  // 3 * pid is even
        o[0] = (uint16_t)(a & 0xffffu);
        o[1] = (uint16_t)((a >> 16) | ((c & 255u) << 8));
        o[2] = (uint16_t)(c >> 8);
    }
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Launch with programmatic stream serialization: the kernel may begin (up to its pdl_wait()) while
// its predecessor in the stream is still running.  GS2D_NO_PDL=1 in the environment turns it off
// (A/B runs).
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args&&... args) {
    static const bool off = [] { const char* e = getenv("GS2D_NO_PDL"); return e && e[0] == '1'; }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = off ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

// launch shape of the tile kernel: one CTA per tile of the band (the experiment build walks the tiles
// with resident CTAs instead and gets the band's height through the flags)
static inline dim3 tile_grid(int tx, int rows) {
#if GS2D_TILE_PERSIST
    const int want = 148 * GS2D_TILE_PERSIST_OCC;
    return dim3(tx * rows < want ? tx * rows : want);
#else
    return dim3(tx, rows);
#endif
}
static inline int tile_flags_for(int flags, int rows) {
#if GS2D_TILE_PERSIST
    return flags | (rows << 8);
#else
    (void)rows;
    return flags;
#endif
}

static int fit_layout(int n, int cap, int img_h, int img_w, gs2d_fit_layout* L) {
    if (n < 0 || cap < 1 || img_h < 1 || img_w < 1) {
        set_error("fit layout: bad sizes (n=%d cap=%d %dx%d)", n, cap, img_w, img_h);
        return GS2D_ERR_ARG;
    }
    const long long tiles = (long long)div_up(img_w, GS2D_TILE) * div_up(img_h, GS2D_TILE);
    size_t o = 0;
    L->off_status = o;      o = align_up(o + 64 * sizeof(int32_t), 256);
    L->off_loss = o;        o = align_up(o + GS2D_FIT_LOSS_SLOTS * sizeof(double), 256);
    L->off_tile_count = o;  o = align_up(o + (size_t)tiles * sizeof(int32_t), 256);
    L->off_tile_off = o;    o = align_up(o + (size_t)(tiles + 1) * sizeof(int32_t), 256);
    L->off_grads = o;       o = align_up(o + (size_t)(n > 0 ? n : 1) * 12 * sizeof(float), 256);
    L->zero_bytes = o;      // everything up to here must be zero before the first step
    L->off_packed = o;      o = align_up(o + (size_t)cap * 3 * sizeof(float4), 256);
    L->total_bytes = o;
    L->tiles = (int)tiles;
    return GS2D_OK;
}

}  // namespace gs2d

using namespace gs2d;

extern "C" {

int gs2d_fit_layout_for(int n, int cap, int img_h, int img_w, gs2d_fit_layout* out) {
    GS2D_REQUIRE(out != nullptr, "fit_layout_for: null out");
    return fit_layout(n, cap, img_h, img_w, out);
}

int gs2d_fit_init(int n, int cap, int img_h, int img_w, void* work, size_t work_bytes, float* adan_state,
                  gs2d_stream_t stream) {
    gs2d_fit_layout L;
    const int rc = fit_layout(n, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(work != nullptr && work_bytes >= L.total_bytes, "fit_init: workspace too small");
    GS2D_REQUIRE(((uintptr_t)work & 255) == 0, "fit_init: workspace must be 256-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    GS2D_CHECK_CUDA(cudaMemsetAsync(work, 0, L.zero_bytes, s));
    if (adan_state && n > 0) GS2D_CHECK_CUDA(cudaMemsetAsync(adan_state, 0, (size_t)n * 8 * 4 * sizeof(float), s));
    return GS2D_OK;
}

static int launch_bin(int n, const float* xyz, const float* chol, const float* feat, const float* opac,
                      const float* rgb_w, int activate,
                      int img_h, int img_w, int row0, int row1, int cap, int step, const gs2d_fit_layout& L, char* w,
                      cudaStream_t s, bool precounted = false, const QuantConst& qc = QuantConst(), bool pdl = false) {
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    const float half_w = 0.5f * (float)(unsigned)img_w, half_h = 0.5f * (float)(unsigned)img_h;
    int32_t* status = (int32_t*)(w + L.off_status);
    int32_t* tile_count = (int32_t*)(w + L.off_tile_count);
    int32_t* tile_off = (int32_t*)(w + L.off_tile_off);
    const int blocks = div_up(n > 0 ? n : 1, kFitThreads);
    if (!precounted) {
        ProfScope ps_(GS2D_STAGE_SCAN_MAP, s);
        k_fit_count<<<blocks, kFitThreads, 0, s>>>(n, xyz, chol, activate, half_w, half_h, tx, ty, row0, row1, cap,
                                                   step, tile_count, tile_off, status, qc);
        GS2D_CHECK_LAUNCH("fit_count");
    }
    {
        ProfScope ps_(GS2D_STAGE_EDGES_PACK, s);
        if (pdl && !precounted)   // render chain: count -> scatter -> tiles, each under its predecessor's tail
            GS2D_CHECK_CUDA(launch_pdl(k_fit_scatter<false>, dim3(blocks), dim3(kFitThreads), s, n, xyz, chol, feat, opac, rgb_w,
                                       activate, half_w, half_h, tx, ty, row0, row1, img_w, img_h, (const int32_t*)tile_off,
                                       tile_count, status, (float4*)(w + L.off_packed), 0, step, qc, Route()));
        else
            k_fit_scatter<false><<<blocks, kFitThreads, 0, s>>>(n, xyz, chol, feat, opac, rgb_w, activate, half_w, half_h, tx, ty,
                                                         row0, row1, img_w, img_h, tile_off, tile_count, status,
                                                         (float4*)(w + L.off_packed), 0, step, qc, Route());
        GS2D_CHECK_LAUNCH("fit_scatter");
    }
    return GS2D_OK;
}

int gs2d_fit_step_band(int n, int cap, int img_h, int img_w, float* xyz_raw, float* chol_raw, float* features,
                       float* adan_state, const void* target, int target_format, void* work, size_t work_bytes,
                       int step, float lr, float beta1, float beta2, float beta3, float eps, float weight_decay,
                       int precounted, int tile_row_begin, int tile_row_end, int phases,
                       const gs2d_grad_exchange* exchange, const gs2d_fit_options* options, float* out_img,
                       float* param_grads_out, int32_t* ids_sorted_out, gs2d_fit_report* report_host,
                       gs2d_stream_t stream) {
    gs2d_fit_layout L;
    int rc = fit_layout(n, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    GS2D_REQUIRE(n > 0 && step >= 1, "fit_step: n and step must be positive");
    GS2D_REQUIRE(work != nullptr && work_bytes >= L.total_bytes, "fit_step: workspace too small");
    GS2D_REQUIRE(target_format == GS2D_TARGET_F32_CHW || target_format == GS2D_TARGET_U8_HWC ||
                     target_format == GS2D_TARGET_I420 || target_format == GS2D_TARGET_TILED, "fit_step: unknown target format");
    GS2D_REQUIRE(target_format != GS2D_TARGET_TILED || ((uintptr_t)target & 15) == 0,
                 "fit_step: a prepared target must be 16-byte aligned");
    GS2D_REQUIRE(target_format != GS2D_TARGET_I420 || (img_h % 2 == 0 && img_w % 2 == 0),
                 "fit_step: an I420 frame needs even width and height");
    GS2D_REQUIRE(xyz_raw && chol_raw && features && adan_state, "fit_step: null pointer");
    GS2D_REQUIRE(0 <= tile_row_begin && tile_row_begin < tile_row_end && tile_row_end <= ty,
                 "fit_step: tile row band out of range");
    GS2D_REQUIRE((phases & ~(GS2D_FIT_PHASE_GRADS | GS2D_FIT_PHASE_UPDATE)) == 0 && phases != 0,
                 "fit_step: bad phase mask");
    const int row0 = tile_row_begin, row1 = tile_row_end;
    const bool banded = !(row0 == 0 && row1 == ty) || phases != (GS2D_FIT_PHASE_GRADS | GS2D_FIT_PHASE_UPDATE);
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)work;
    const float half_w = 0.5f * (float)(unsigned)img_w, half_h = 0.5f * (float)(unsigned)img_h;
    int32_t* status = (int32_t*)(w + L.off_status);
    double* loss = (double*)(w + L.off_loss);
    float* grads = (float*)(w + L.off_grads);
    const float* const* peer_grads = nullptr;
    int num_peers = 0;
    float* grads_zero = nullptr;
    const float* mc_grads = nullptr;
    Route rt = Route();
    float* const* push_peers = nullptr;
    int push_per = 1;
    int shard_begin = 0, shard_end = n;
    if (exchange) {
        GS2D_REQUIRE(exchange->accumulate != nullptr, "fit_step: exchange without an accumulate buffer");
        GS2D_REQUIRE(exchange->num_peers >= 0 && (exchange->num_peers == 0 || exchange->peers != nullptr),
                     "fit_step: bad peer list");
        grads = exchange->accumulate;
        peer_grads = exchange->peers;
        num_peers = exchange->num_peers;
        grads_zero = exchange->zero;
        mc_grads = exchange->multicast;
        if (mc_grads) { peer_grads = nullptr; num_peers = 0; }
        if (exchange->push_per > 0) {
            GS2D_REQUIRE(exchange->shard_end > exchange->shard_begin && exchange->peers != nullptr && !mc_grads,
                         "fit_step: a pushed reduction needs a sharded update and the peers' buffers");
            push_peers = (float* const*)exchange->peers;
            push_per = exchange->push_per;
        }
        if (exchange->shard_end > exchange->shard_begin) {
            GS2D_REQUIRE(exchange->shard_begin >= 0 && exchange->shard_end <= n, "fit_step: shard out of range");
            GS2D_REQUIRE(num_peers > 0 && exchange->xyz_peers && exchange->status_peers && (grads_zero || push_peers) &&
                             ((options && options->params8) || (exchange->chol_peers && exchange->feat_peers)),
                         "fit_step: a sharded update needs peers, their parameter pointers and (pulled reduction) a spare buffer");
            shard_begin = exchange->shard_begin;
            shard_end = exchange->shard_end;
            rt.xyz = exchange->xyz_peers;
            rt.chol = exchange->chol_peers;
            rt.feat = exchange->feat_peers;
            rt.status = exchange->status_peers;
            rt.R = num_peers;
            if (exchange->work_peers) {
                GS2D_REQUIRE(exchange->band_edges != nullptr, "fit_step: remote scatter needs the band edges");
                rt.work = (char* const*)exchange->work_peers;
                rt.off_tile_count = L.off_tile_count;
                rt.off_packed = L.off_packed;
                rt.band_edges = exchange->band_edges;
            }
        }
    }
    const bool sharded = rt.R > 0;
    // slot mode drops the pairs whose pixel rectangle in the tile is empty (scatter_records); a call that asks
    // for the rendering order (ids_sorted_out: parity tests) and scatters itself keeps them, so that the list
    // is the reference's.  GS2D_NO_CULL=1: A/B runs.
    static const bool no_cull = [] { const char* e = getenv("GS2D_NO_CULL"); return e && e[0] == '1'; }();
    const int cull_bit = no_cull ? 0 : 2;
    float* rgb_w = nullptr; float* adan_w = nullptr; float* rgb_w_grad_out = nullptr;
    QuantConst src = QuantConst();   // row strides of the parameter tensors (packed rows: 8 / 8 / 8)
    int adan_step = step, adan_fresh = step == 1, skip_update = 0, slots = 0, analytic = 0, all_per_tile = 0;
    if (options) {
        slots = options->slots != 0;
        GS2D_REQUIRE(!slots || (long long)cap >= (long long)L.tiles * GS2D_MAX_PER_TILE,
                     "fit_step: slot mode needs cap >= tiles * 256");
        GS2D_REQUIRE(options->rgb_w == nullptr || options->adan_state_w != nullptr,
                     "fit_step: rgb_w without its optimizer state");
        GS2D_REQUIRE(options->adan_step >= 0, "fit_step: negative adan_step");
        rgb_w = options->rgb_w; adan_w = options->adan_state_w; rgb_w_grad_out = options->rgb_w_grad_out;
        if (options->adan_step > 0) { adan_step = options->adan_step; adan_fresh = options->adan_fresh != 0; }
        skip_update = options->skip_update != 0;
        analytic = options->analytic_grad != 0;
        all_per_tile = options->all_per_tile != 0;
        if (options->params8) {
            GS2D_REQUIRE(((uintptr_t)options->params8 & 15) == 0, "fit_step: params8 must be 16-byte aligned");
            GS2D_REQUIRE(xyz_raw == options->params8 && chol_raw == options->params8 + 2 && features == options->params8 + 5,
                         "fit_step: with params8 the three parameter pointers must be its columns 0 / 2 / 5");
            src = packed8_rows();
        }
        GS2D_REQUIRE(!(all_per_tile && slots), "fit_step: all_per_tile needs tile ranges (a slot holds 256 records)");
    }
    GS2D_REQUIRE(!sharded || (slots && rgb_w == nullptr && !skip_update),
                 "fit_step: a sharded update needs tile slots (and has no prune / add control)");
    if (phases & GS2D_FIT_PHASE_GRADS) {
        GS2D_REQUIRE(target != nullptr, "fit_step: null target");
        if (slots) {
            // precounted == 0: nothing scattered for these parameters yet (first step, or they were
            // modified).  A sharded update never scatters: the band's records are written here, every
            // step, from the parameters the ranks stored into this rank's tensors (the fill levels
            // are zero already: the previous tile kernel left them so)
            if (!precounted || (sharded && !rt.work)) {   // (remote scatter: the peers' update kernels filled the slots)
                ProfScope ps_(GS2D_STAGE_EDGES_PACK, s);
                if (!precounted) {
                    k_fit_reset_slots<<<div_up(L.tiles, kFitThreads), kFitThreads, 0, s>>>(
                        L.tiles, (int32_t*)(w + L.off_tile_count), status);
                    GS2D_CHECK_LAUNCH("fit_reset_slots");
                }
                if (rt.R > 0) {
                    // launched programmatically: its blocks are resident when the barrier kernel before it (the
                    // previous step's second barrier) lets go; it reads nothing before its wait
                    GS2D_CHECK_CUDA(launch_pdl(k_fit_scatter<true>, dim3(div_up(n, kFitThreads)), dim3(kFitThreads), s,
                        n, (const float*)xyz_raw, (const float*)chol_raw, (const float*)features, (const float*)nullptr,
                        (const float*)rgb_w, 1, half_w, half_h, tx, ty, row0, row1, img_w, img_h, (const int32_t*)nullptr,
                        (int32_t*)(w + L.off_tile_count), status, (float4*)(w + L.off_packed),
                        1 | (ids_sorted_out ? 0 : cull_bit), step, src, rt));
                } else {
                    k_fit_scatter<false><<<div_up(n, kFitThreads), kFitThreads, 0, s>>>(
                        n, xyz_raw, chol_raw, features, nullptr, rgb_w, 1, half_w, half_h, tx, ty, row0, row1, img_w, img_h,
                        nullptr, (int32_t*)(w + L.off_tile_count), status, (float4*)(w + L.off_packed),
                        1 | (ids_sorted_out ? 0 : cull_bit), step, src, rt);
                }
                GS2D_CHECK_LAUNCH("fit_scatter_slots");
            }
        } else {
            rc = launch_bin(n, xyz_raw, chol_raw, features, nullptr, rgb_w, 1, img_h, img_w, row0, row1, cap, step, L, w,
                            s, precounted != 0, src);
            if (rc != GS2D_OK) return rc;
        }
        const int32_t* tile_off_arg = slots ? nullptr : (const int32_t*)(w + L.off_tile_off);
        const float grad_scale = (float)(2.0 / (3.0 * (double)img_h * (double)img_w));   // the WHOLE frame's mean
        const int tile_flags = tile_flags_for(all_per_tile ? kTileAllPerTile : 0,
                                              row1 - row0);
        ProfScope ps_(GS2D_STAGE_FIT_TILES, s);
        const dim3 grid = tile_grid(tx, row1 - row0);
        // where the band's loss also goes: the pad of gradient row 0, or (pushed reduction, where row 0 belongs to
        // rank 0 alone) one of two words past the rows, by step parity (gs2d.h, push_per)
        float* loss_pad = !banded ? nullptr : push_peers ? grads + (size_t)n * 12 + 16 * (step & 1) : grads + 9;
        auto launch = [&](auto kernel) {
            return launch_pdl(kernel, grid, dim3(kTileThreads), s, img_w, img_h, tx, row0, tile_off_arg,
                              (int32_t*)(w + L.off_tile_count), (const float4*)(w + L.off_packed), target, grad_scale, grads,
                              loss + (step % GS2D_FIT_LOSS_SLOTS), status, out_img, 1, 1, (int32_t*)nullptr, ids_sorted_out,
                              step, tile_flags, push_peers, push_per, loss_pad);
        };
        cudaError_t le;
        if (push_peers) {   // (slots only: checked above)
            le = target_format == GS2D_TARGET_F32_CHW  ? launch(k_fit_tiles<0, 0, false, true>)
                 : target_format == GS2D_TARGET_U8_HWC ? launch(k_fit_tiles<0, 1, false, true>)
                 : target_format == GS2D_TARGET_I420   ? launch(k_fit_tiles<0, 2, false, true>)
                                                       : launch(k_fit_tiles<0, 3, false, true>);
        } else if (target_format == GS2D_TARGET_F32_CHW)
            le = all_per_tile ? launch(k_fit_tiles<0, 0, true>) : launch(k_fit_tiles<0, 0, false>);
        else if (target_format == GS2D_TARGET_U8_HWC)
            le = all_per_tile ? launch(k_fit_tiles<0, 1, true>) : launch(k_fit_tiles<0, 1, false>);
        else if (target_format == GS2D_TARGET_I420)
            le = all_per_tile ? launch(k_fit_tiles<0, 2, true>) : launch(k_fit_tiles<0, 2, false>);
        else
            le = all_per_tile ? launch(k_fit_tiles<0, 3, true>) : launch(k_fit_tiles<0, 3, false>);
        GS2D_CHECK_CUDA(le);
        GS2D_CHECK_LAUNCH("fit_tiles");
    }
    const bool own_barrier = exchange && exchange->barrier_pads &&
                             phases == (GS2D_FIT_PHASE_GRADS | GS2D_FIT_PHASE_UPDATE);
    if (own_barrier) {
        GS2D_REQUIRE(exchange->num_peers > 0 && exchange->num_peers <= 32 && exchange->rank >= 0 &&
                         exchange->rank < exchange->num_peers, "fit_step: bad barrier description");
        GS2D_CHECK_CUDA(launch_pdl(k_rank_barrier, dim3(1), dim3(32), s, exchange->barrier_pads, exchange->num_peers,
                                   exchange->rank, exchange->barrier_epoch));
        GS2D_CHECK_LAUNCH("rank_barrier");
    }
    if (phases & GS2D_FIT_PHASE_UPDATE) {
        AdanHyper h;
        h.b1 = beta1; h.b2 = beta2; h.b3 = beta3;
        h.step_size = (float)((double)lr / (1.0 - pow((double)beta1, adan_step)));
        h.step_size_diff = (float)((double)lr * (double)beta2 / (1.0 - pow((double)beta2, adan_step)));
        h.inv_bc3_sqrt = (float)(1.0 / sqrt(1.0 - pow((double)beta3, adan_step)));
        h.eps = eps;
        h.decay_div = 1.f + lr * weight_decay;
        h.first = adan_fresh;
        h.offdiag = analytic ? 1.f : 2.f;
        ProfScope ps_(GS2D_STAGE_ADAN, s);
        static const bool no_quad = [] { const char* e = getenv("GS2D_NO_QUAD"); return e && e[0] == '1'; }();   // A/B runs
        GS2D_REQUIRE(!push_peers || (src.ld_xyz == 8 && !no_quad), "fit_step: a pushed reduction needs the packed parameter rows");
        if (src.ld_xyz == 8 && !no_quad)
            GS2D_CHECK_CUDA(launch_pdl((exchange || rt.R > 0) ? k_fit_update_quad<true, kUpdateLanes> : k_fit_update_quad<false, kUpdateLanes>,
                dim3(div_up((long long)kUpdateLanes * (shard_end - shard_begin), kFitThreads)),
                dim3(kFitThreads), s, n, xyz_raw, adan_state, grads, half_w, half_h, tx, ty, h, status,
                loss + (step % GS2D_FIT_LOSS_SLOTS), loss + ((step + 1) % GS2D_FIT_LOSS_SLOTS), param_grads_out, cap,
                step + 1, (int32_t*)(w + L.off_tile_count), (int32_t*)(w + L.off_tile_off), report_host, row0, row1,
                (banded ? 1 : 0) | (push_peers ? 2 : 0), peer_grads, num_peers, grads_zero, rgb_w, adan_w, rgb_w_grad_out,
                skip_update, mc_grads, slots ? (1 | cull_bit) : 0, (float4*)(w + L.off_packed), img_w, img_h, shard_begin,
                shard_end, rt));
        else
        GS2D_CHECK_CUDA(launch_pdl(src.ld_xyz == 8 ? k_fit_update<true> : k_fit_update<false>,
            dim3(div_up(shard_end - shard_begin, kFitThreads)), dim3(kFitThreads), s,
            n, xyz_raw, chol_raw, features, adan_state, grads, half_w, half_h, tx, ty, h, status,
            loss + (step % GS2D_FIT_LOSS_SLOTS), loss + ((step + 1) % GS2D_FIT_LOSS_SLOTS), param_grads_out, cap,
            step + 1, (int32_t*)(w + L.off_tile_count), (int32_t*)(w + L.off_tile_off), report_host, row0, row1,
            banded ? 1 : 0, peer_grads, num_peers, grads_zero, rgb_w, adan_w, rgb_w_grad_out, skip_update,
            mc_grads, slots ? (1 | cull_bit) : 0, (float4*)(w + L.off_packed), img_w, img_h, shard_begin, shard_end, rt));
        GS2D_CHECK_LAUNCH("fit_update");
    }
    if (own_barrier && sharded) {   // the peers' stores into this rank's parameters precede its next scatter
        GS2D_CHECK_CUDA(launch_pdl(k_rank_barrier, dim3(1), dim3(32), s, exchange->barrier_pads, exchange->num_peers,
                                   exchange->rank, exchange->barrier_epoch + 1u));
        GS2D_CHECK_LAUNCH("rank_barrier");
    }
    return GS2D_OK;
}

size_t gs2d_fit_target_bytes(int img_h, int img_w) {
    if (img_h < 1 || img_w < 1) return 0;
    return (size_t)div_up(img_w, GS2D_TILE) * div_up(img_h, GS2D_TILE) * GS2D_TILE_PIX * 3 * sizeof(float);
}

int gs2d_fit_target_prepare(int img_h, int img_w, const void* target, int target_format, float* prepared,
                            gs2d_stream_t stream) {
    GS2D_REQUIRE(img_h > 0 && img_w > 0 && target && prepared, "fit_target_prepare: bad arguments");
    GS2D_REQUIRE(((uintptr_t)prepared & 15) == 0, "fit_target_prepare: output must be 16-byte aligned");
    GS2D_REQUIRE(target_format != GS2D_TARGET_I420 || (img_h % 2 == 0 && img_w % 2 == 0),
                 "fit_target_prepare: an I420 frame needs even width and height");
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 grid(div_up(img_w, GS2D_TILE), div_up(img_h, GS2D_TILE));
    ProfScope ps_(GS2D_STAGE_OTHER, s);
    if (target_format == GS2D_TARGET_F32_CHW)
        k_target_tiled<0><<<grid, kTileThreads, 0, s>>>(img_w, img_h, grid.x, target, prepared);
    else if (target_format == GS2D_TARGET_U8_HWC)
        k_target_tiled<1><<<grid, kTileThreads, 0, s>>>(img_w, img_h, grid.x, target, prepared);
    else if (target_format == GS2D_TARGET_I420)
        k_target_tiled<2><<<grid, kTileThreads, 0, s>>>(img_w, img_h, grid.x, target, prepared);
    else
        GS2D_REQUIRE(false, "fit_target_prepare: unknown target format");
    GS2D_CHECK_LAUNCH("target_tiled");
    return GS2D_OK;
}

int gs2d_fit_step(int n, int cap, int img_h, int img_w, float* xyz_raw, float* chol_raw, float* features,
                  float* adan_state, const void* target, int target_format, void* work, size_t work_bytes, int step,
                  float lr, float beta1, float beta2, float beta3, float eps, float weight_decay, int precounted,
                  float* out_img, float* param_grads_out, int32_t* ids_sorted_out, gs2d_fit_report* report_host,
                  gs2d_stream_t stream) {
    GS2D_REQUIRE(img_h > 0, "fit_step: bad sizes");
    return gs2d_fit_step_band(n, cap, img_h, img_w, xyz_raw, chol_raw, features, adan_state, target, target_format,
                              work, work_bytes, step, lr, beta1, beta2, beta3, eps, weight_decay, precounted, 0,
                              div_up(img_h, GS2D_TILE), GS2D_FIT_PHASE_GRADS | GS2D_FIT_PHASE_UPDATE, nullptr, nullptr,
                              out_img, param_grads_out, ids_sorted_out, report_host, stream);
}

static int quant_const(const gs2d_quantized_frame* q, QuantConst& qc) {
    GS2D_REQUIRE(q != nullptr && q->xyz_half && q->chol_code && q->color_index && q->codebooks,
                 "quantized frame: null pointer");
    GS2D_REQUIRE(((uintptr_t)q->xyz_half & 3) == 0, "quantized frame: xyz_half must be 4-byte aligned");
    for (int c = 0; c < 3; ++c) { qc.scale[c] = q->chol_scale[c]; qc.beta[c] = q->chol_beta[c]; }
    qc.codebooks = q->codebooks;
    return GS2D_OK;
}

int gs2d_render_quantized(int n, int cap, int img_h, int img_w, const gs2d_quantized_frame* q, void* work,
                          size_t work_bytes, float* out_img, int out_chw, int out_clamp, int32_t* final_idx,
                          int32_t* ids_sorted_out, gs2d_stream_t stream) {
    gs2d_fit_layout L;
    int rc = fit_layout(n, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(n > 0, "render_quantized: n must be positive");
    GS2D_REQUIRE(work != nullptr && work_bytes >= L.total_bytes, "render_quantized: workspace too small");
    GS2D_REQUIRE(out_img != nullptr, "render_quantized: null output");
    QuantConst qc = QuantConst();
    rc = quant_const(q, qc);
    if (rc != GS2D_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)work;
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    rc = launch_bin(n, (const float*)q->xyz_half, (const float*)q->chol_code, (const float*)q->color_index, nullptr,
                    nullptr, 2, img_h, img_w, 0, ty, cap, 0, L, w, s, false, qc, true);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        GS2D_CHECK_CUDA(launch_pdl(k_fit_tiles<1, 0, false>, tile_grid(tx, ty), dim3(kTileThreads), s,
            img_w, img_h, tx, 0, (const int32_t*)(w + L.off_tile_off), (int32_t*)(w + L.off_tile_count),
            (const float4*)(w + L.off_packed), (const void*)nullptr, 0.f, (float*)nullptr, (double*)nullptr,
            (int32_t*)(w + L.off_status), out_img, out_chw, out_clamp, final_idx, ids_sorted_out, 0,
            tile_flags_for(0, ty), (float* const*)nullptr, 1, (float*)nullptr));
        GS2D_CHECK_LAUNCH("render_tiles");
    }
    return GS2D_OK;
}

int gs2d_dequantize(int n, const gs2d_quantized_frame* q, float* xyz_raw, float* chol_raw, float* features,
                    gs2d_stream_t stream) {
    GS2D_REQUIRE(n > 0 && xyz_raw && chol_raw && features, "dequantize: bad arguments");
    QuantConst qc = QuantConst();
    const int rc = quant_const(q, qc);
    if (rc != GS2D_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope ps_(GS2D_STAGE_OTHER, s);
    k_dequantize<<<div_up(n, kFitThreads), kFitThreads, 0, s>>>(n, q->xyz_half, q->chol_code, q->color_index, qc,
                                                               xyz_raw, chol_raw, features);
    GS2D_CHECK_LAUNCH("dequantize");
    return GS2D_OK;
}

int gs2d_i420_to_rgb(int img_h, int img_w, const uint8_t* yuv, uint8_t* rgb, gs2d_stream_t stream) {
    GS2D_REQUIRE(img_h > 0 && img_w > 0 && img_h % 2 == 0 && img_w % 2 == 0, "i420_to_rgb: width and height must be even");
    GS2D_REQUIRE(yuv && rgb, "i420_to_rgb: null pointer");
    GS2D_REQUIRE((((uintptr_t)yuv | (uintptr_t)rgb) & 1) == 0, "i420_to_rgb: buffers must be 2-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const int blocks2x2 = (img_h / 2) * (img_w / 2);
    ProfScope ps_(GS2D_STAGE_OTHER, s);
    k_i420_to_rgb<<<div_up(blocks2x2, kFitThreads), kFitThreads, 0, s>>>(img_h, img_w, yuv, rgb);
    GS2D_CHECK_LAUNCH("i420_to_rgb");
    return GS2D_OK;
}

size_t gs2d_opstate_bytes(int cap, int img_h, int img_w) {
    gs2d_fit_layout L;
    if (fit_layout(1, cap, img_h, img_w, &L) != GS2D_OK) return 0;
    return L.total_bytes;
}

int gs2d_rasterize_sum_fwd_ranked(int n, const float* xys, const int32_t* radii, const float* conics,
                                  const float* colors, const float* opacities, const float* background, int img_h,
                                  int img_w, int cap, void* state, size_t state_bytes, float* out_img,
                                  int32_t* final_idx, int32_t* mail_host, int mail_seq, gs2d_stream_t stream) {
    gs2d_fit_layout L;
    int rc = fit_layout(1, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(n > 0, "rasterize_sum_fwd_ranked: n must be positive");
    GS2D_REQUIRE(state != nullptr && state_bytes >= L.total_bytes, "rasterize_sum_fwd_ranked: state buffer too small");
    GS2D_REQUIRE(((uintptr_t)state & 255) == 0, "rasterize_sum_fwd_ranked: state buffer must be 256-byte aligned");
    GS2D_REQUIRE(xys && radii && conics && colors && opacities && out_img, "rasterize_sum_fwd_ranked: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)state;
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    {
        ProfScope ps_(GS2D_STAGE_MEMSET, s);
        GS2D_CHECK_CUDA(cudaMemsetAsync(w, 0, L.zero_bytes, s));
    }
    QuantConst qc = QuantConst();
    qc.radii = radii;
    qc.mail_host = mail_host;
    qc.mail_seq = mail_seq;
    rc = launch_bin(n, xys, conics, colors, opacities, nullptr, 3, img_h, img_w, 0, ty, cap, 0, L, w, s, false, qc, true);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        GS2D_CHECK_CUDA(launch_pdl(k_fit_tiles<1, 0, false>, tile_grid(tx, ty), dim3(kTileThreads), s,
            img_w, img_h, tx, 0, (const int32_t*)(w + L.off_tile_off), (int32_t*)(w + L.off_tile_count),
            (const float4*)(w + L.off_packed), (const void*)background, 0.f, (float*)nullptr, (double*)nullptr,
            (int32_t*)(w + L.off_status), out_img, 0, 0, final_idx, (int32_t*)nullptr, 0, tile_flags_for(0, ty),
            (float* const*)nullptr, 1, (float*)nullptr));
        GS2D_CHECK_LAUNCH("raster_fwd_ranked");
    }
    return GS2D_OK;
}

int gs2d_opstate_ranges(int cap, int img_h, int img_w, size_t* off_tile_off, size_t* off_packed, size_t* off_status) {
    gs2d_fit_layout L;
    const int rc = fit_layout(1, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    if (off_tile_off) *off_tile_off = L.off_tile_off;
    if (off_packed) *off_packed = L.off_packed;
    if (off_status) *off_status = L.off_status;
    return GS2D_OK;
}

int gs2d_render_chol(int n, int cap, int img_h, int img_w, const float* means_or_xyz_raw, const float* chol,
                     const float* colors, const float* opacities, int activate, void* work, size_t work_bytes,
                     float* out_img, int out_chw, int out_clamp, int32_t* final_idx, int32_t* ids_sorted_out,
                     gs2d_stream_t stream) {
    gs2d_fit_layout L;
    int rc = fit_layout(n, cap, img_h, img_w, &L);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(n > 0, "render_chol: n must be positive");
    GS2D_REQUIRE(work != nullptr && work_bytes >= L.total_bytes, "render_chol: workspace too small");
    GS2D_REQUIRE(means_or_xyz_raw && chol && colors && out_img, "render_chol: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)work;
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    const int all_flag = (activate & GS2D_RENDER_ALL_PER_TILE) ? kTileAllPerTile : 0;
    QuantConst src = QuantConst();
    if (activate & GS2D_RENDER_PACKED8) {
        GS2D_REQUIRE(((uintptr_t)means_or_xyz_raw & 15) == 0 && chol == means_or_xyz_raw + 2,
                     "render_chol: packed rows need a 16-byte aligned [n][8] buffer and chol = its column 2");
        src.ld_xyz = src.ld_chol = 8;
        src.ld_feat = colors == means_or_xyz_raw + 5 ? 8 : 0;   // colours may be a tensor of their own (features * rgb_W)
    }
    rc = launch_bin(n, means_or_xyz_raw, chol, colors, opacities, nullptr, (activate & GS2D_RENDER_ACTIVATE) ? 1 : 0, img_h,
                    img_w, 0, ty, cap, 0, L, w, s, false, src, true);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        GS2D_CHECK_CUDA(launch_pdl(all_flag ? k_fit_tiles<1, 0, true> : k_fit_tiles<1, 0, false>, tile_grid(tx, ty),
            dim3(kTileThreads), s,
            img_w, img_h, tx, 0, (const int32_t*)(w + L.off_tile_off), (int32_t*)(w + L.off_tile_count),
            (const float4*)(w + L.off_packed), (const void*)nullptr, 0.f, (float*)nullptr, (double*)nullptr,
            (int32_t*)(w + L.off_status), out_img, out_chw, out_clamp, final_idx, ids_sorted_out, 0,
            tile_flags_for(all_flag, ty), (float* const*)nullptr, 1, (float*)nullptr));
        GS2D_CHECK_LAUNCH("render_tiles");
    }
    return GS2D_OK;
}

}  // extern "C"
