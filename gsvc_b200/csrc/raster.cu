// raster.cu -- accumulated-sum rasteriser, forward and backward.
// Reference: rasterize_forward_sum forward.cu:512-627 (host bindings.cu:400-469),
//            rasterize_backward_sum_kernel backward.cu:696-862 (host bindings.cu:706-779).
//
// Forward: one CTA per 16x16 tile, one thread per pixel.  The tile's batch (at most the
// first 256 sorted entries -- the reference renders only batch 0) is one
// contiguous run of 48-byte packed records, staged into shared memory with 16-byte
// cp.async.  Each warp owns two pixel rows; 32 lanes test 32 records' row ranges at once
// and the warp only evaluates the records whose conservative rectangle touches its rows
// (warp-cooperative early-out; rejected pixels could not pass the alpha test, so the
// result is unchanged).
//
// Backward: one CTA per tile stages v_out / final_idx of its 256 pixels and its records in
// shared memory once; eight lanes per (tile, Gaussian) pair walk the pair's pixel rectangle
// column by column, a transposed 7-shuffle reduction leaves one total per lane, and the pair
// issues a single coalesced reduction (8 lanes, consecutive floats) plus one for opacity --
// one atomic transaction per Gaussian per tile, where the reference issues 9 atomicAdd per
// warp per Gaussian.  (First version: one warp per pair reading v_out from global memory,
// 76 us at 1080p / 50k against 35 us now; the N-channel variant below still has that shape.)
#include "gs2d_internal.h"
#include "gs2d_math.cuh"
#include "gs2d_stage.cuh"

namespace gs2d {

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GS2D_TILE_PIX)
k_raster_fwd(int W, int H, int tx, const int32_t* __restrict__ tile_bins, const float4* __restrict__ packed,
             float* __restrict__ out_img, float* __restrict__ final_Ts, int32_t* __restrict__ final_idx) {
    __shared__ float4 sq[GS2D_MAX_PER_TILE * 3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = blockIdx.y * tx + blockIdx.x;
    const int2 range = reinterpret_cast<const int2*>(tile_bins)[t];
    const int lo = range.x;
    const int nb = max(0, min(range.y - range.x, GS2D_MAX_PER_TILE));

    const float4* src = packed + 3 * (size_t)lo;
    for (int i = tid; i < nb * 3; i += GS2D_TILE_PIX) cp_async16(&sq[i], src + i);
    cp_async_wait_all();
    __syncthreads();

    const int lx = tid & 15, ly = tid >> 4;   // warp w owns rows 2w, 2w+1 of the tile
    const int px_i = blockIdx.x * GS2D_TILE + lx, py_i = blockIdx.y * GS2D_TILE + ly;
    const bool inside = px_i < W && py_i < H;
    const float px = (float)px_i, py = (float)py_i;
    const int row_lo = 2 * warp, row_hi = 2 * warp + 1;

    float3 pix = make_float3(0.f, 0.f, 0.f);
    int cur_idx = 0;
    for (int c = 0; c < nb; c += 32) {
        const int k_lane = c + lane;
        bool touch = false;
        if (k_lane < nb) {
            const uint32_t rect = __float_as_uint(sq[3 * k_lane + 1].z);
            const int y0 = (rect >> 16) & 255, y1 = rect >> 24;
            touch = (y1 >= row_lo) && (y0 <= row_hi) && (y1 >= y0);   // (empty rectangle: y0 = 1, y1 = 0)
        }
        unsigned m = __ballot_sync(0xffffffffu, touch);
        while (m) {
            const int k = c + __ffs(m) - 1;
            m &= m - 1;
            const float4 q0 = sq[3 * k];       // x, y, conic.x, conic.y
            const float4 q1 = sq[3 * k + 1];   // conic.z, opacity, rect, id
            const float dx = subf(q0.x, px), dy = subf(q0.y, py);
            const float sigma = gauss_sigma(dx, dy, q0.z, q0.w, q1.x);
            const float alpha = minf(mulf(q1.y, gauss_vis(sigma)), 1.f);
            if (inside && gauss_pass(sigma, alpha)) {
                const float4 q2 = sq[3 * k + 2];   // r, g, b, tile
                pix.x = fmaf_(alpha, q2.x, pix.x);
                pix.y = fmaf_(alpha, q2.y, pix.y);
                pix.z = fmaf_(alpha, q2.z, pix.z);
                cur_idx = lo + k;
            }
        }
    }
    if (inside) {
        const size_t p = (size_t)py_i * W + px_i;
        out_img[3 * p] = pix.x; out_img[3 * p + 1] = pix.y; out_img[3 * p + 2] = pix.z;
        if (final_idx) final_idx[p] = cur_idx;
        if (final_Ts) final_Ts[p] = 1.f;   // constant in the sum variant (forward.cu:557, 618)
    }
}

// ------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------
// Sum 8 per-lane values over the warp with 9 shuffles: after the call, the total of v[j]
// is held by the lanes whose (lane >> 2) == j.
__device__ __forceinline__ float warp_reduce8_transposed(float (&v)[8], int lane) {
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    float w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = b4 ? v[k] : v[4 + k];
        const float keep = b4 ? v[4 + k] : v[k];
        w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float u[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float send = b3 ? w[k] : w[2 + k];
        const float keep = b3 ? w[2 + k] : w[k];
        u[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    const float send = b2 ? u[0] : u[1];
    const float keep = b2 ? u[1] : u[0];
    float r = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    r += __shfl_xor_sync(0xffffffffu, r, 2);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

constexpr int kBwdThreads = 256;   // N-channel backward: one warp per (tile, Gaussian) pair

// ------------------------------------------------------------------------------------------
// backward, tile-cooperative
// ------------------------------------------------------------------------------------------
// Sum 8 per-lane values over each group of 8 consecutive lanes with 7 shuffles: lane l of a
// group ends up with the group's total of v[l].
__device__ __forceinline__ float group8_reduce8(float (&v)[8], int lane) {
    const bool b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
    float w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = b2 ? v[k] : v[4 + k];
        const float keep = b2 ? v[4 + k] : v[k];
        w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    float u[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float send = b1 ? w[k] : w[2 + k];
        const float keep = b1 ? w[2 + k] : w[k];
        u[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    const float send = b0 ? u[0] : u[1];
    const float keep = b0 ? u[1] : u[0];
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

// One CTA of two warps per tile.  v_out and final_idx of the tile's 256 pixels are read ONCE,
// into shared memory (the warp-per-pair kernel re-read them for every Gaussian of the tile: 40 MB
// of DRAM traffic and ~300 MB of L2 reads at 1080p / 50k); the tile's first 256 sorted records
//  arrive by cp.async.  Then, as in the fused fit kernel (fit.cu), eight lanes per
// (tile, Gaussian) pair: a lane owns one column of the pair's rectangle (two when it is wider
// than 8) and walks its rows, accumulating the moments of v_sigma over dy; one transposed
// 7-shuffle reduction per pass and one reduction instruction for four pairs' 8 floats (+ one
// for the opacities).  Rows past a pair's rectangle need no mask: inside the tile the rectangle
// guarantees that the alpha test fails, past the tile the staged v_out rows are zero.
constexpr int kBwdTileThreads = 64;
// RANKED = false: records sorted in memory (onesweep path; tile_bins rows of (lo, hi), conics as given);
// RANKED = true:  records as the count / scatter binning of fit.cu leaves them -- a tile's range is
//                 [tile_off[t], tile_off[t+1]), in arbitrary order, conic diagonals stored halved --
//                 and the batch is put into ascending-id order while it is staged (gs2d_stage.cuh),
//                 the same order the forward (k_fit_tiles<1>) rendered and numbered final_idx in.
template <bool RANKED>
__global__ void __launch_bounds__(kBwdTileThreads)
k_raster_bwd_tiles(int W, int H, int tx, const int32_t* __restrict__ tile_bins, const float4* __restrict__ packed,
                   const int32_t* __restrict__ final_idx, const float* __restrict__ v_out,
                   float* __restrict__ grads, const int32_t* __restrict__ status) {
    __shared__ float4 sq_all[GS2D_MAX_PER_TILE * 3];
    __shared__ float4 sv[2 * GS2D_TILE_PIX];   // {v_out.rgb, final_idx} of rows 0..15, then 16 rows of zeros
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = blockIdx.y * tx + blockIdx.x;
    int lo, cnt;
    if (RANKED) {
        if (status[1] != 0) return;   // the forward overflowed its capacity: nothing was rendered
        lo = tile_bins[t];
        cnt = tile_bins[t + 1] - lo;
    } else {
        const int2 range = reinterpret_cast<const int2*>(tile_bins)[t];
        lo = range.x;
        cnt = range.y - range.x;
    }
    const int nb = max(0, min(cnt, GS2D_MAX_PER_TILE));
    if (nb == 0) return;
    const float4* src = packed + 3 * (size_t)lo;
    float4* sq = sq_all;
    const bool per_warp = RANKED && cnt <= 32;
    if (!RANKED) {
        for (int i = tid; i < nb * 3; i += kBwdTileThreads) cp_async16(&sq[i], src + i);
    } else if (per_warp) {
        sq = sq_all + warp * (32 * 3);
        float4 w0 = make_float4(0.f, 0.f, 0.f, 0.f), w1 = w0, w2 = w0;
        if (lane < cnt) { w0 = __ldg(src + 3 * lane); w1 = __ldg(src + 3 * lane + 1); w2 = __ldg(src + 3 * lane + 2); }
        stage_tile_warp(w0, w1, w2, cnt, sq, lane);
    } else {
        // ids go through the zero rows of sv (re-zeroed below, before anyone reads them as v_out)
        stage_tile_sorted<kBwdTileThreads>(src, cnt, sq, reinterpret_cast<int*>(sv + GS2D_TILE_PIX), tid);
    }
    const int ox = blockIdx.x * GS2D_TILE, oy = blockIdx.y * GS2D_TILE;
#pragma unroll
    for (int j = 0; j < GS2D_TILE_PIX / kBwdTileThreads; ++j) {
        const int p = tid + j * kBwdTileThreads;
        const int px_i = ox + (p & 15), py_i = oy + (p >> 4);
        float4 v = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));   // outside the image: nothing passes the index test
        if (px_i < W && py_i < H) {
            const size_t pid = (size_t)py_i * W + px_i;
            v = make_float4(__ldg(v_out + 3 * pid), __ldg(v_out + 3 * pid + 1), __ldg(v_out + 3 * pid + 2),
                            __int_as_float(__ldg(final_idx + pid)));
        }
        sv[p] = v;
        sv[GS2D_TILE_PIX + p] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
    }
    if (!RANKED) cp_async_wait_all();
    __syncthreads();

    const int sub = lane >> 3, l8 = lane & 7;
    for (int r0 = warp * 4; r0 < nb; r0 += (kBwdTileThreads / 32) * 4) {
        const int k = min(r0 + sub, nb - 1);
        const float4 q1 = sq[3 * k + 1];
        const uint32_t rect = __float_as_uint(q1.z);
        const int rx0 = rect & 255, rx1 = (rect >> 8) & 255, ry0 = (rect >> 16) & 255, ry1 = rect >> 24;
        const bool live = (r0 + sub < nb) && ry1 >= ry0;
        GS2D_ASSERT(!live || (rx0 <= rx1 && rx1 < GS2D_TILE && ry1 < GS2D_TILE));
        const int h = live ? ry1 - ry0 + 1 : 0;
        const int ncol = live ? ((rx1 - rx0) >> 3) + 1 : 0;   // column blocks of 8: 1 or 2
        int max_iters = h * ncol;
        max_iters = max(max_iters, __shfl_xor_sync(0xffffffffu, max_iters, 8));
        max_iters = max(max_iters, __shfl_xor_sync(0xffffffffu, max_iters, 16));
        if (max_iters == 0) continue;
        const float4 q0 = sq[3 * k], q2 = sq[3 * k + 2];
        const float gx = q0.x, gy = q0.y, cy = q0.w, opac = q1.y;
        // halving / doubling is exact: sigma keeps the reference's roundings either way (fit.cu)
        const float cxh = RANKED ? q0.z : 0.5f * q0.z, czh = RANKED ? q1.x : 0.5f * q1.x;
        const float cx = RANKED ? 2.f * q0.z : q0.z, cz = RANKED ? 2.f * q1.x : q1.x;
        const int gidx = lo + k;                        // sorted index of this pair (backward.cu:784-787)
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float acc_op = 0.f;
        for (int cb0 = 0; cb0 < 2; ++cb0) {
            int max_h = (cb0 < ncol) ? h : 0;
            max_h = max(max_h, __shfl_xor_sync(0xffffffffu, max_h, 8));
            max_h = max(max_h, __shfl_xor_sync(0xffffffffu, max_h, 16));
            if (max_h == 0) break;
            const int lxp = rx0 + l8 + 8 * cb0;
            const bool col_ok = (cb0 < ncol) && lxp <= rx1;
            const int lxs = min(lxp, GS2D_TILE - 1);
            const float dx = subf(gx, (float)(ox + lxp));
            const float t1 = mulf(cxh, dx), t3 = mulf(cy, dx);
            float S0 = 0.f, S1 = 0.f, S2 = 0.f, cr = 0.f, cg = 0.f, cbl = 0.f, op = 0.f;
            float dy0 = subf(gy, (float)(oy + ry0));   // exact, and so is stepping it by whole rows
            const float4* svc = sv + ry0 * GS2D_TILE + lxs;
            // (a lane outside its pair's columns is not masked per pixel: its sums are dropped after the walk)
#pragma unroll 1
            for (int ry = 0; ry < max_h; ry += 2) {
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const float dy = dy0 - (float)u;
                    const float sg = fmaf_(dy, t3, fmaf_(dx, t1, mulf(mulf(czh, dy), dy)));
                    const float vis = gauss_vis(sg);
                    const float al = minf(mulf(opac, vis), 1.f);
                    const float4 vo = svc[(ry + u) * GS2D_TILE];
                    const bool ok = gidx <= __float_as_int(vo.w) && gauss_pass(sg, al);
                    const float visk = ok ? vis : 0.f, alk = ok ? al : 0.f;
                    const float va = visk * (q2.x * vo.x + q2.y * vo.y + q2.z * vo.z);   // vis * v_alpha  (backward.cu:821-825)
                    const float v_sigma = -opac * va;                                      // :827
                    const float tt = v_sigma * dy;
                    op += va;                                                              // v_opacity :833
                    S0 += v_sigma;
                    S1 += tt;
                    S2 = fmaf(tt, dy, S2);
                    cr = fmaf(alk, vo.x, cr);                                              // v_rgb :817
                    cg = fmaf(alk, vo.y, cg);
                    cbl = fmaf(alk, vo.z, cbl);
                }
                dy0 -= 2.f;
            }
            if (!col_ok) { S0 = 0.f; S1 = 0.f; S2 = 0.f; cr = 0.f; cg = 0.f; cbl = 0.f; op = 0.f; }
            acc_op += op;
            acc[0] += cx * dx * S0 + cy * S1;          // v_xy     :831-832
            acc[1] += cy * dx * S0 + cz * S1;
            acc[2] += 0.5f * dx * dx * S0;             // v_conic  :828-830
            acc[3] += 0.5f * dx * S1;
            acc[4] += 0.5f * S2;
            acc[5] += cr; acc[6] += cg; acc[7] += cbl;
        }
        const float tot = group8_reduce8(acc, lane);
        acc_op += __shfl_xor_sync(0xffffffffu, acc_op, 4);
        acc_op += __shfl_xor_sync(0xffffffffu, acc_op, 2);
        acc_op += __shfl_xor_sync(0xffffffffu, acc_op, 1);
        float* row = grads + (size_t)__float_as_int(q1.w) * 12;
        if (live && tot != 0.f) atomicAdd(row + l8, tot);   // 4 pairs x 8 consecutive floats
        if (live && l8 == 0 && acc_op != 0.f) atomicAdd(row + 8, acc_op);
    }
}

// ------------------------------------------------------------------------------------------
// N-channel / untruncated variant (no bound counterpart in the reference: its
// nd_rasterize_forward_sum forward.cu:630-710 and nd_rasterize_backward_sum_kernel
// backward.cu:864-978 exist but are never launched).
//
// Same tile / record layout as above, but (1) the colours of a batch are staged by Gaussian id
// into shared memory as [record][channel] rows next to the geometry records, (2) a tile is
// rendered in batches of 256 records up to `max_tile` (INT_MAX = every Gaussian of the tile, as
// the reference's N-channel kernel does; 256 = the 3-channel kernel's truncation),
// (3) colour gradients go to their own [n][C] tensor.  The pixel test and the accumulation order
// are the 3-channel kernel's (the reference's N-channel SASS has the identical sigma / EX2 /
// FFMA sequence), so with C = 3 and max_tile = 256 the image is bit-identical to k_raster_fwd.
// final_idx keeps the 3-channel meaning (last sorted index that passed, 0 when none).
// ------------------------------------------------------------------------------------------
template <int CMAX>
__global__ void __launch_bounds__(GS2D_TILE_PIX)
k_raster_nd_fwd(int W, int H, int tx, int C, int max_tile, const int32_t* __restrict__ tile_bins,
                const float4* __restrict__ packed, const float* __restrict__ colors, float* __restrict__ out_img,
                int32_t* __restrict__ final_idx) {
    extern __shared__ float4 nd_smem[];
    float4* sq = nd_smem;                                                    // [256][3]
    float* scol = reinterpret_cast<float*>(nd_smem + GS2D_MAX_PER_TILE * 3);   // [256][C]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = blockIdx.y * tx + blockIdx.x;
    const int2 range = reinterpret_cast<const int2*>(tile_bins)[t];
    const int lo = range.x;
    const int cnt = max(0, min(range.y - range.x, max_tile));

    const int lx = tid & 15, ly = tid >> 4;
    const int px_i = blockIdx.x * GS2D_TILE + lx, py_i = blockIdx.y * GS2D_TILE + ly;
    const bool inside = px_i < W && py_i < H;
    const float px = (float)px_i, py = (float)py_i;
    const int row_lo = 2 * warp, row_hi = 2 * warp + 1;

    float acc[CMAX];
#pragma unroll
    for (int c = 0; c < CMAX; ++c) acc[c] = 0.f;
    int cur_idx = 0;
    for (int base = 0; base < cnt; base += GS2D_MAX_PER_TILE) {
        const int nb = min(GS2D_MAX_PER_TILE, cnt - base);
        if (base) __syncthreads();   // everyone is done with the previous batch
        const float4* src = packed + 3 * (size_t)(lo + base);
        for (int i = tid; i < nb * 3; i += GS2D_TILE_PIX) cp_async16(&sq[i], src + i);
        cp_async_wait_all();
        __syncthreads();
        for (int i = tid; i < nb * C; i += GS2D_TILE_PIX) {
            const int k = i / C;
            const int g = __float_as_int(sq[3 * k + 1].w);
            scol[i] = colors[(size_t)g * C + (i - k * C)];
        }
        __syncthreads();
        for (int c0 = 0; c0 < nb; c0 += 32) {
            const int k_lane = c0 + lane;
            bool touch = false;
            if (k_lane < nb) {
                const uint32_t rect = __float_as_uint(sq[3 * k_lane + 1].z);
                const int y0 = (rect >> 16) & 255, y1 = rect >> 24;
                touch = (y1 >= row_lo) && (y0 <= row_hi) && (y1 >= y0);   // (empty rectangle: y0 = 1, y1 = 0)
            }
            unsigned m = __ballot_sync(0xffffffffu, touch);
            while (m) {
                const int k = c0 + __ffs(m) - 1;
                m &= m - 1;
                const float4 q0 = sq[3 * k];
                const float4 q1 = sq[3 * k + 1];
                const float dx = subf(q0.x, px), dy = subf(q0.y, py);
                const float sigma = gauss_sigma(dx, dy, q0.z, q0.w, q1.x);
                const float alpha = minf(mulf(q1.y, gauss_vis(sigma)), 1.f);
                if (inside && gauss_pass(sigma, alpha)) {
                    const float* col = scol + k * C;
#pragma unroll
                    for (int c = 0; c < CMAX; ++c)
                        if (c < C) acc[c] = fmaf_(alpha, col[c], acc[c]);
                    cur_idx = lo + base + k;
                }
            }
        }
    }
    if (inside) {
        const size_t p = (size_t)py_i * W + px_i;
#pragma unroll
        for (int c = 0; c < CMAX; ++c)
            if (c < C) out_img[p * C + c] = acc[c];
        if (final_idx) final_idx[p] = cur_idx;
    }
}

// one warp per sorted (tile, Gaussian) pair sweeping the pair's rectangle; the C colour sums are reduced in
// groups of eight with the same transposed butterfly.
template <int CMAX>
__global__ void __launch_bounds__(kBwdThreads)
k_raster_nd_bwd(const int32_t* __restrict__ n_dev, int n_static, int W, int tx, int C, int max_tile,
                const int32_t* __restrict__ tile_bins, const float4* __restrict__ packed,
                const float* __restrict__ colors, const int32_t* __restrict__ final_idx,
                const float* __restrict__ v_out, float* __restrict__ grads, float* __restrict__ v_colors) {
    const int n = n_dev ? *n_dev : n_static;
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (kBwdThreads / 32) + (threadIdx.x >> 5);
    if (i >= n) return;
    const float4 q1 = packed[3 * (size_t)i + 1];
    const uint32_t rect = __float_as_uint(q1.z);
    const int rx0 = rect & 255, rx1 = (rect >> 8) & 255, ry0 = (rect >> 16) & 255, ry1 = rect >> 24;
    if (ry1 < ry0) return;
    const int t = (int)__float_as_uint(packed[3 * (size_t)i + 2].w);
    if (i - tile_bins[2 * t] >= max_tile) return;
    const float4 q0 = packed[3 * (size_t)i];
    const float gx = q0.x, gy = q0.y, cx = q0.z, cy = q0.w, cz = q1.x, opac = q1.y;
    const int g = __float_as_int(q1.w);
    float col[CMAX], vcol[CMAX];
#pragma unroll
    for (int c = 0; c < CMAX; ++c) {
        col[c] = c < C ? colors[(size_t)g * C + c] : 0.f;
        vcol[c] = 0.f;
    }
    const int ox = (t % tx) * GS2D_TILE, oy = (t / tx) * GS2D_TILE;
    const int w = rx1 - rx0 + 1, area = w * (ry1 - ry0 + 1);
    const float inv_w = 1.f / (float)w;

    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // v_xy 2 | v_conic 3 | v_opacity | - | -
    for (int p = lane; p < area; p += 32) {
        const int ry = (int)(((float)p + 0.5f) * inv_w);
        const int px_i = ox + rx0 + (p - ry * w), py_i = oy + ry0 + ry;
        const size_t pid = (size_t)py_i * W + px_i;
        if (i > final_idx[pid]) continue;
        const float dx = subf(gx, (float)px_i), dy = subf(gy, (float)py_i);
        const float sigma = gauss_sigma(dx, dy, cx, cy, cz);
        const float vis = gauss_vis(sigma);
        const float alpha = minf(mulf(opac, vis), 1.f);
        if (!gauss_pass(sigma, alpha)) continue;
        const float* vo = v_out + pid * C;
        float v_alpha = 0.f;
#pragma unroll
        for (int c = 0; c < CMAX; ++c) {
            if (c < C) {
                const float v = vo[c];
                v_alpha += col[c] * v;
                vcol[c] += alpha * v;
            }
        }
        const float v_sigma = -opac * vis * v_alpha;
        acc[0] += v_sigma * (cx * dx + cy * dy);
        acc[1] += v_sigma * (cy * dx + cz * dy);
        const float hs = 0.5f * v_sigma;
        acc[2] += hs * dx * dx;
        acc[3] += hs * dx * dy;
        acc[4] += hs * dy * dy;
        acc[5] += vis * v_alpha;
    }
    const float tot = warp_reduce8_transposed(acc, lane);
    float* row = grads + (size_t)g * 12;
    if ((lane & 3) == 0 && tot != 0.f) {
        const int j = lane >> 2;
        if (j < 5) atomicAdd(row + j, tot);
        else if (j == 5) atomicAdd(row + 8, tot);
    }
#pragma unroll
    for (int c8 = 0; c8 < CMAX; c8 += 8) {
        if (c8 < C) {
            float part[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) part[k] = c8 + k < CMAX ? vcol[c8 + k] : 0.f;
            const float ct = warp_reduce8_transposed(part, lane);
            const int c = c8 + (lane >> 2);
            if ((lane & 3) == 0 && c < C && ct != 0.f) atomicAdd(v_colors + (size_t)g * C + c, ct);
        }
    }
}

constexpr int kMaxChannels = 32;
static size_t nd_smem_bytes(int C) { return (size_t)GS2D_MAX_PER_TILE * (3 * sizeof(float4) + (size_t)C * sizeof(float)); }

template <int CMAX>
static cudaError_t launch_nd_fwd(dim3 grid, int W, int H, int tx, int C, int max_tile, const int32_t* tile_bins,
                                 const float4* packed, const float* colors, float* out_img, int32_t* final_idx,
                                 cudaStream_t s) {
    const size_t smem = nd_smem_bytes(C);
    if (smem > 48 * 1024) {
        const cudaError_t e = cudaFuncSetAttribute(k_raster_nd_fwd<CMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   (int)nd_smem_bytes(CMAX));
        if (e != cudaSuccess) return e;
    }
    k_raster_nd_fwd<CMAX><<<grid, GS2D_TILE_PIX, smem, s>>>(W, H, tx, C, max_tile, tile_bins, packed, colors, out_img,
                                                           final_idx);
    return cudaSuccess;
}

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace gs2d

using namespace gs2d;

extern "C" {

int gs2d_rasterize_sum_fwd(int n, const float* xys, const float* depths, const int32_t* radii, const float* conics,
                           const int32_t* num_tiles_hit, const float* colors, const float* opacities, int img_h,
                           int img_w, int cap, void* state, size_t state_bytes, float* out_img, int32_t* final_idx,
                           int32_t* status_host, gs2d_event_t scan_done, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && cap > 0, "rasterize_sum_fwd: bad sizes");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    StateOffsets so;
    int rc = compute_state_offsets(n, cap, tx, ty, &so);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(state_bytes >= so.pub.total_bytes, "rasterize_sum_fwd: state buffer too small");
    GS2D_REQUIRE(((uintptr_t)state & 255) == 0, "rasterize_sum_fwd: state buffer must be 256-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    char* st = (char*)state;
    rc = launch_bin_and_pack(n, xys, depths, radii, num_tiles_hit, conics, colors, opacities, img_h, img_w, cap, so,
                             st, status_host, (cudaEvent_t)scan_done, s);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        k_raster_fwd<<<dim3(tx, ty), GS2D_TILE_PIX, 0, s>>>(img_w, img_h, tx, (const int32_t*)(st + so.pub.off_tile_bins),
                                                        (const float4*)(st + so.pub.off_packed), out_img, nullptr,
                                                        final_idx);
    }
    GS2D_CHECK_LAUNCH("raster_fwd");
    return GS2D_OK;
}

size_t gs2d_raster_binned_scratch_bytes(int num_intersects) {
    return align256((size_t)(num_intersects > 0 ? num_intersects : 1) * 3 * sizeof(float4));
}

int gs2d_rasterize_sum_fwd_binned(int n, int num_intersects, const int32_t* ids_sorted, const int32_t* tile_bins,
                                  const float* xys, const float* conics, const float* colors, const float* opacities,
                                  int img_h, int img_w, float* out_img, float* final_Ts, int32_t* final_idx,
                                  void* scratch, size_t scratch_bytes, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && num_intersects >= 0 && img_h > 0 && img_w > 0, "rasterize_sum_fwd_binned: bad sizes");
    GS2D_REQUIRE(scratch_bytes >= gs2d_raster_binned_scratch_bytes(num_intersects),
                 "rasterize_sum_fwd_binned: scratch too small");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = launch_pack_binned(num_intersects, ids_sorted, tile_bins, tx * ty, xys, conics, colors, opacities, img_h,
                                img_w, (float4*)scratch, s);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        k_raster_fwd<<<dim3(tx, ty), GS2D_TILE_PIX, 0, s>>>(img_w, img_h, tx, tile_bins, (const float4*)scratch, out_img,
                                                        final_Ts, final_idx);
    }
    GS2D_CHECK_LAUNCH("raster_fwd_binned");
    return GS2D_OK;
}

int gs2d_rasterize_sum_bwd(int n, int img_h, int img_w, int cap, const void* state, size_t state_bytes,
                           const int32_t* final_idx, const float* v_out, float* grads, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && cap > 0, "rasterize_sum_bwd: bad sizes");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    StateOffsets so;
    const int rc = compute_state_offsets(n, cap, tx, ty, &so);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(state_bytes >= so.pub.total_bytes, "rasterize_sum_bwd: state buffer too small");
    cudaStream_t s = (cudaStream_t)stream;
    const char* st = (const char*)state;
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_MEMSET, s);
        GS2D_CHECK_CUDA(cudaMemsetAsync(grads, 0, (size_t)n * 12 * sizeof(float), s));
    }
    {
        ProfScope ps_(GS2D_STAGE_RASTER_BWD, s);
        k_raster_bwd_tiles<false><<<dim3(tx, ty), kBwdTileThreads, 0, s>>>(
        img_w, img_h, tx, (const int32_t*)(st + so.pub.off_tile_bins), (const float4*)(st + so.pub.off_packed),
        final_idx, v_out, grads, nullptr);
    }
    GS2D_CHECK_LAUNCH("raster_bwd");
    return GS2D_OK;
}

int gs2d_rasterize_sum_bwd_ranked(int n, int img_h, int img_w, int cap, const void* state, size_t state_bytes,
                                  const int32_t* final_idx, const float* v_out, float* grads, gs2d_stream_t stream) {
    GS2D_REQUIRE(n > 0 && img_h > 0 && img_w > 0 && cap > 0, "rasterize_sum_bwd_ranked: bad sizes");
    size_t off_tile_off = 0, off_packed = 0, off_status = 0;
    const int rc = gs2d_opstate_ranges(cap, img_h, img_w, &off_tile_off, &off_packed, &off_status);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(state != nullptr && state_bytes >= gs2d_opstate_bytes(cap, img_h, img_w),
                 "rasterize_sum_bwd_ranked: state buffer too small");
    GS2D_REQUIRE(final_idx && v_out && grads, "rasterize_sum_bwd_ranked: null pointer");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    cudaStream_t s = (cudaStream_t)stream;
    const char* st = (const char*)state;
    {
        ProfScope ps_(GS2D_STAGE_MEMSET, s);
        GS2D_CHECK_CUDA(cudaMemsetAsync(grads, 0, (size_t)n * 12 * sizeof(float), s));
    }
    {
        ProfScope ps_(GS2D_STAGE_RASTER_BWD, s);
        k_raster_bwd_tiles<true><<<dim3(tx, ty), kBwdTileThreads, 0, s>>>(
            img_w, img_h, tx, (const int32_t*)(st + off_tile_off), (const float4*)(st + off_packed), final_idx, v_out,
            grads, (const int32_t*)(st + off_status));
    }
    GS2D_CHECK_LAUNCH("raster_bwd_ranked");
    return GS2D_OK;
}

int gs2d_rasterize_sum_bwd_binned(int n, int num_intersects, const int32_t* ids_sorted, const int32_t* tile_bins,
                                  const float* xys, const float* conics, const float* colors, const float* opacities,
                                  int img_h, int img_w, const int32_t* final_idx, const float* v_out, float* grads,
                                  void* scratch, size_t scratch_bytes, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && num_intersects >= 0 && img_h > 0 && img_w > 0, "rasterize_sum_bwd_binned: bad sizes");
    GS2D_REQUIRE(scratch_bytes >= gs2d_raster_binned_scratch_bytes(num_intersects),
                 "rasterize_sum_bwd_binned: scratch too small");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) return GS2D_OK;
    GS2D_CHECK_CUDA(cudaMemsetAsync(grads, 0, (size_t)n * 12 * sizeof(float), s));
    if (num_intersects == 0) return GS2D_OK;
    // entries that belong to no tile range keep an all-zero record: opacity 0 never passes the alpha test
    GS2D_CHECK_CUDA(cudaMemsetAsync(scratch, 0, gs2d_raster_binned_scratch_bytes(num_intersects), s));
    int rc = launch_pack_binned(num_intersects, ids_sorted, tile_bins, tx * ty, xys, conics, colors, opacities, img_h,
                                img_w, (float4*)scratch, s);
    if (rc != GS2D_OK) return rc;
    {
        ProfScope ps_(GS2D_STAGE_RASTER_BWD, s);
        k_raster_bwd_tiles<false><<<dim3(tx, ty), kBwdTileThreads, 0, s>>>(img_w, img_h, tx, tile_bins,
                                                                          (const float4*)scratch, final_idx, v_out,
                                                                          grads, nullptr);
    }
    GS2D_CHECK_LAUNCH("raster_bwd_binned");
    return GS2D_OK;
}

int gs2d_rasterize_sum_nd_fwd(int n, int channels, const float* xys, const float* depths, const int32_t* radii,
                              const float* conics, const int32_t* num_tiles_hit, const float* colors,
                              const float* opacities, int img_h, int img_w, int cap, int max_per_tile, void* state,
                              size_t state_bytes, float* out_img, int32_t* final_idx, int32_t* status_host,
                              gs2d_event_t scan_done, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && cap > 0 && max_per_tile >= 0, "rasterize_sum_nd_fwd: bad sizes");
    GS2D_REQUIRE(channels >= 1 && channels <= kMaxChannels, "rasterize_sum_nd_fwd: 1..32 channels are supported");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    StateOffsets so;
    int rc = compute_state_offsets(n, cap, tx, ty, &so);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(state_bytes >= so.pub.total_bytes, "rasterize_sum_nd_fwd: state buffer too small");
    GS2D_REQUIRE(((uintptr_t)state & 255) == 0, "rasterize_sum_nd_fwd: state buffer must be 256-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    char* st = (char*)state;
    rc = launch_bin_and_pack(n, xys, depths, radii, num_tiles_hit, conics, nullptr, opacities, img_h, img_w, cap, so,
                             st, status_host, (cudaEvent_t)scan_done, s);
    if (rc != GS2D_OK) return rc;
    const int max_tile = max_per_tile > 0 ? max_per_tile : 0x7fffffff;
    const int32_t* bins = (const int32_t*)(st + so.pub.off_tile_bins);
    const float4* packed = (const float4*)(st + so.pub.off_packed);
    const dim3 grid(tx, ty);
    {
        ProfScope ps_(GS2D_STAGE_RASTER_FWD, s);
        cudaError_t e;
        if (channels <= 4) e = launch_nd_fwd<4>(grid, img_w, img_h, tx, channels, max_tile, bins, packed, colors, out_img, final_idx, s);
        else if (channels <= 8) e = launch_nd_fwd<8>(grid, img_w, img_h, tx, channels, max_tile, bins, packed, colors, out_img, final_idx, s);
        else if (channels <= 16) e = launch_nd_fwd<16>(grid, img_w, img_h, tx, channels, max_tile, bins, packed, colors, out_img, final_idx, s);
        else e = launch_nd_fwd<32>(grid, img_w, img_h, tx, channels, max_tile, bins, packed, colors, out_img, final_idx, s);
        GS2D_CHECK_CUDA(e);
    }
    GS2D_CHECK_LAUNCH("raster_nd_fwd");
    return GS2D_OK;
}

int gs2d_rasterize_sum_nd_bwd(int n, int channels, int img_h, int img_w, int cap, int max_per_tile, const void* state,
                              size_t state_bytes, const float* colors, const int32_t* final_idx, const float* v_out,
                              float* grads, float* v_colors, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && cap > 0 && max_per_tile >= 0, "rasterize_sum_nd_bwd: bad sizes");
    GS2D_REQUIRE(channels >= 1 && channels <= kMaxChannels, "rasterize_sum_nd_bwd: 1..32 channels are supported");
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    StateOffsets so;
    const int rc = compute_state_offsets(n, cap, tx, ty, &so);
    if (rc != GS2D_OK) return rc;
    GS2D_REQUIRE(state_bytes >= so.pub.total_bytes, "rasterize_sum_nd_bwd: state buffer too small");
    cudaStream_t s = (cudaStream_t)stream;
    const char* st = (const char*)state;
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_MEMSET, s);
        GS2D_CHECK_CUDA(cudaMemsetAsync(grads, 0, (size_t)n * 12 * sizeof(float), s));
        GS2D_CHECK_CUDA(cudaMemsetAsync(v_colors, 0, (size_t)n * channels * sizeof(float), s));
    }
    const int max_tile = max_per_tile > 0 ? max_per_tile : 0x7fffffff;
    const int32_t* nd = (const int32_t*)(st + so.pub.off_status) + 3;
    const int32_t* bins = (const int32_t*)(st + so.pub.off_tile_bins);
    const float4* packed = (const float4*)(st + so.pub.off_packed);
    const int blocks = div_up(cap, kBwdThreads / 32);
    {
        ProfScope ps_(GS2D_STAGE_RASTER_BWD, s);
        if (channels <= 4) k_raster_nd_bwd<4><<<blocks, kBwdThreads, 0, s>>>(nd, 0, img_w, tx, channels, max_tile, bins, packed, colors, final_idx, v_out, grads, v_colors);
        else if (channels <= 8) k_raster_nd_bwd<8><<<blocks, kBwdThreads, 0, s>>>(nd, 0, img_w, tx, channels, max_tile, bins, packed, colors, final_idx, v_out, grads, v_colors);
        else if (channels <= 16) k_raster_nd_bwd<16><<<blocks, kBwdThreads, 0, s>>>(nd, 0, img_w, tx, channels, max_tile, bins, packed, colors, final_idx, v_out, grads, v_colors);
        else k_raster_nd_bwd<32><<<blocks, kBwdThreads, 0, s>>>(nd, 0, img_w, tx, channels, max_tile, bins, packed, colors, final_idx, v_out, grads, v_colors);
    }
    GS2D_CHECK_LAUNCH("raster_nd_bwd");
    return GS2D_OK;
}

}  // extern "C"
