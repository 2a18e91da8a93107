/*
 * ref_launch.cu -- C-ABI launcher around the REFERENCE's own CUDA kernels.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/gs2d_oracle.c header).  The kernels themselves
 * are compiled by oracle/build_ref.sh straight from /root/reference/gsplat/gsplat/cuda/csrc
 * (forward.cu, foward2d.cu, backward.cu, backward2d.cu -- no reference source is copied
 * into this repo); this file only declares them through the reference's own headers and
 * launches them with the grid/block shapes and zero-initialised outputs that the
 * reference's host code uses (bindings.cu, cited per function).  bindings.cu itself needs
 * torch and pybind11 and is not built; its one kernel on the path
 * (compute_cov2d_bounds_kernel, bindings.cu:21-39) is a 10-line wrapper around the device
 * function helpers.cuh:45-68, which is called here directly.
 *
 * All pointers are device pointers.  Launches go to `stream` (0 = legacy default stream,
 * which is what the reference uses).
 */
#include <cstdint>
#include <cuda_runtime.h>

#include "forward.cuh"
#include "backward.cuh"
#include "forward2d.cuh"
#include "backward2d.cuh"
#include "helpers.cuh"

#define REF_THREADS 256 /* config.h:4 N_THREADS */

static inline int nblocks(int n) { return (n + REF_THREADS - 1) / REF_THREADS; }

__global__ void ref_cov2d_bounds_wrapper(unsigned n, const float *covs2d, float *conics, float *radii) {
    unsigned row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    float3 conic; float radius;
    float3 cov2d{covs2d[3 * row], covs2d[3 * row + 1], covs2d[3 * row + 2]};
    compute_cov2d_bounds(cov2d, conic, radius);
    conics[3 * row] = conic.x; conics[3 * row + 1] = conic.y; conics[3 * row + 2] = conic.z;
    radii[row] = radius;
}

extern "C" {

int ref_last_error() { return (int)cudaGetLastError(); }

/* bindings.cu:41-61 */
int ref_compute_cov2d_bounds(int n, const float *covs2d, float *conics, float *radii, cudaStream_t s) {
    cudaMemsetAsync(conics, 0, sizeof(float) * 3 * n, s);
    cudaMemsetAsync(radii, 0, sizeof(float) * n, s);
    ref_cov2d_bounds_wrapper<<<nblocks(n), REF_THREADS, 0, s>>>(n, covs2d, conics, radii);
    return (int)cudaGetLastError();
}

/* bindings.cu:787-839 */
int ref_project_gaussians_2d_forward(int n, const float *means2d, const float *L, int H, int W, int tx, int ty,
                                     float clip_thresh, float *xys, float *depths, int *radii, float *conics,
                                     int *num_tiles_hit, cudaStream_t s) {
    dim3 img(W, H, 1), tb(tx, ty, 1);
    cudaMemsetAsync(xys, 0, sizeof(float) * 2 * n, s); cudaMemsetAsync(depths, 0, sizeof(float) * n, s);
    cudaMemsetAsync(radii, 0, sizeof(int) * n, s); cudaMemsetAsync(conics, 0, sizeof(float) * 3 * n, s);
    cudaMemsetAsync(num_tiles_hit, 0, sizeof(int) * n, s);
    project_gaussians_2d_forward_kernel<<<nblocks(n), REF_THREADS, 0, s>>>(
        n, (const float2 *)means2d, (const float3 *)L, img, tb, clip_thresh, (float2 *)xys, depths, radii,
        (float3 *)conics, num_tiles_hit);
    return (int)cudaGetLastError();
}

/* bindings.cu:841-899 */
int ref_project_gaussians_2d_scale_rot_forward(int n, const float *means2d, const float *scales, const float *rot,
                                               int H, int W, int tx, int ty, float clip_thresh, float *xys,
                                               float *depths, int *radii, float *conics, int *num_tiles_hit,
                                               cudaStream_t s) {
    dim3 img(W, H, 1), tb(tx, ty, 1);
    cudaMemsetAsync(xys, 0, sizeof(float) * 2 * n, s); cudaMemsetAsync(depths, 0, sizeof(float) * n, s);
    cudaMemsetAsync(radii, 0, sizeof(int) * n, s); cudaMemsetAsync(conics, 0, sizeof(float) * 3 * n, s);
    cudaMemsetAsync(num_tiles_hit, 0, sizeof(int) * n, s);
    project_gaussians_2d_scale_rot_forward_kernel<<<nblocks(n), REF_THREADS, 0, s>>>(
        n, (const float2 *)means2d, (const float2 *)scales, rot, img, tb, clip_thresh, (float2 *)xys, depths, radii,
        (float3 *)conics, num_tiles_hit);
    return (int)cudaGetLastError();
}

/* bindings.cu:902-949 */
int ref_project_gaussians_2d_backward(int n, const float *means2d, const float *L, int H, int W, const int *radii,
                                      const float *conics, const float *v_xy, const float *v_depth,
                                      const float *v_conic, float *v_cov2d, float *v_mean2d, float *v_L,
                                      cudaStream_t s) {
    dim3 img(W, H, 1);
    cudaMemsetAsync(v_cov2d, 0, sizeof(float) * 3 * n, s); cudaMemsetAsync(v_mean2d, 0, sizeof(float) * 2 * n, s);
    cudaMemsetAsync(v_L, 0, sizeof(float) * 3 * n, s);
    project_gaussians_2d_backward_kernel<<<nblocks(n), REF_THREADS, 0, s>>>(
        n, (const float2 *)means2d, (const float3 *)L, img, radii, (const float3 *)conics, (const float2 *)v_xy,
        v_depth, (const float3 *)v_conic, (float3 *)v_cov2d, (float2 *)v_mean2d, (float3 *)v_L);
    return (int)cudaGetLastError();
}

/* bindings.cu:951-1004 */
int ref_project_gaussians_2d_scale_rot_backward(int n, const float *means2d, const float *scales, const float *rot,
                                                int H, int W, const int *radii, const float *conics,
                                                const float *v_xy, const float *v_depth, const float *v_conic,
                                                float *v_cov2d, float *v_mean2d, float *v_scale, float *v_rot,
                                                cudaStream_t s) {
    dim3 img(W, H, 1);
    cudaMemsetAsync(v_cov2d, 0, sizeof(float) * 3 * n, s); cudaMemsetAsync(v_mean2d, 0, sizeof(float) * 2 * n, s);
    cudaMemsetAsync(v_scale, 0, sizeof(float) * 2 * n, s); cudaMemsetAsync(v_rot, 0, sizeof(float) * n, s);
    project_gaussians_2d_scale_rot_backward_kernel<<<nblocks(n), REF_THREADS, 0, s>>>(
        n, (const float2 *)means2d, (const float2 *)scales, rot, img, radii, (const float3 *)conics,
        (const float2 *)v_xy, v_depth, (const float3 *)v_conic, (float3 *)v_cov2d, (float2 *)v_mean2d,
        (float2 *)v_scale, v_rot);
    return (int)cudaGetLastError();
}

/* bindings.cu:274-313 */
int ref_map_gaussian_to_intersects(int n, int num_intersects, const float *xys, const float *depths,
                                   const int *radii, const int *cum_tiles_hit, int tx, int ty, int64_t *isect_ids,
                                   int *gaussian_ids, cudaStream_t s) {
    dim3 tb(tx, ty, 1);
    cudaMemsetAsync(isect_ids, 0, sizeof(int64_t) * num_intersects, s);
    cudaMemsetAsync(gaussian_ids, 0, sizeof(int) * num_intersects, s);
    map_gaussian_to_intersects<<<nblocks(n), REF_THREADS, 0, s>>>(n, (const float2 *)xys, depths, radii,
                                                                 cum_tiles_hit, tb, isect_ids, gaussian_ids);
    return (int)cudaGetLastError();
}

/* bindings.cu:315-330.  `rows` = number of (lo,hi) rows the caller allocated; the reference
 * allocates num_intersects rows -- callers here pass max(I, tiles) to stay in bounds. */
int ref_get_tile_bin_edges(int num_intersects, const int64_t *isect_ids_sorted, int *tile_bins, int rows,
                           cudaStream_t s) {
    cudaMemsetAsync(tile_bins, 0, sizeof(int) * 2 * rows, s);
    get_tile_bin_edges<<<nblocks(num_intersects), REF_THREADS, 0, s>>>(num_intersects, isect_ids_sorted,
                                                                        (int2 *)tile_bins);
    return (int)cudaGetLastError();
}

/* bindings.cu:400-469.  `background` is a device pointer to 3 floats (bound as a reference
 * parameter in the kernel signature; unused by the kernel body). */
int ref_rasterize_sum_forward(int H, int W, const int *gaussian_ids_sorted, const int *tile_bins, const float *xys,
                              const float *conics, const float *colors, const float *opacities,
                              const float *background, float *out_img, float *final_Ts, int *final_idx,
                              cudaStream_t s) {
    dim3 tb((W + 15) / 16, (H + 15) / 16, 1), block(16, 16, 1), img(W, H, 1);
    cudaMemsetAsync(out_img, 0, sizeof(float) * 3 * H * W, s);
    cudaMemsetAsync(final_Ts, 0, sizeof(float) * H * W, s);
    cudaMemsetAsync(final_idx, 0, sizeof(int) * H * W, s);
    rasterize_forward_sum<<<tb, block, 0, s>>>(tb, img, gaussian_ids_sorted, (const int2 *)tile_bins,
                                               (const float2 *)xys, (const float3 *)conics, (const float3 *)colors,
                                               opacities, final_Ts, final_idx, (float3 *)out_img,
                                               *(const float3 *)background);
    return (int)cudaGetLastError();
}

/* bindings.cu:706-779 */
int ref_rasterize_sum_backward(int n, int H, int W, const int *gaussian_ids_sorted, const int *tile_bins,
                               const float *xys, const float *conics, const float *colors, const float *opacities,
                               const float *background, const float *final_Ts, const int *final_idx,
                               const float *v_output, const float *v_output_alpha, float *v_xy, float *v_conic,
                               float *v_colors, float *v_opacity, cudaStream_t s) {
    dim3 tb((W + 15) / 16, (H + 15) / 16, 1), block(16, 16, 1), img(W, H, 1);
    cudaMemsetAsync(v_xy, 0, sizeof(float) * 2 * n, s); cudaMemsetAsync(v_conic, 0, sizeof(float) * 3 * n, s);
    cudaMemsetAsync(v_colors, 0, sizeof(float) * 3 * n, s); cudaMemsetAsync(v_opacity, 0, sizeof(float) * n, s);
    rasterize_backward_sum_kernel<<<tb, block, 0, s>>>(
        tb, img, gaussian_ids_sorted, (const int2 *)tile_bins, (const float2 *)xys, (const float3 *)conics,
        (const float3 *)colors, opacities, *(const float3 *)background, final_Ts, final_idx,
        (const float3 *)v_output, v_output_alpha, (float2 *)v_xy, (float3 *)v_conic, (float3 *)v_colors, v_opacity);
    return (int)cudaGetLastError();
}

/* The reference's N-channel sum kernels (forward.cu:630-710, backward.cu:864-978).  The reference
 * has NO host code for them (its rasterize_sum.py:147-150, 218-221 names functions that ext.cpp
 * never binds); they are launched here the way bindings.cu launches their
 * alpha-compositing siblings (nd_rasterize_forward_tensor bindings.cu:475-549): one 16x16 block per
 * tile, zero-initialised outputs (the forward accumulates into out_img with +=). */
int ref_nd_rasterize_sum_forward(int H, int W, int channels, const int *gaussian_ids_sorted, const int *tile_bins,
                                 const float *xys, const float *conics, const float *colors, const float *opacities,
                                 const float *background, float *out_img, float *final_Ts, int *final_idx,
                                 cudaStream_t s) {
    dim3 tb((W + 15) / 16, (H + 15) / 16, 1), block(16, 16, 1), img(W, H, 1);
    cudaMemsetAsync(out_img, 0, sizeof(float) * channels * H * W, s);
    cudaMemsetAsync(final_Ts, 0, sizeof(float) * H * W, s);
    cudaMemsetAsync(final_idx, 0, sizeof(int) * H * W, s);
    nd_rasterize_forward_sum<<<tb, block, 0, s>>>(tb, img, (unsigned)channels, gaussian_ids_sorted,
                                                  (const int2 *)tile_bins, (const float2 *)xys,
                                                  (const float3 *)conics, colors, opacities, final_Ts, final_idx,
                                                  out_img, background);
    return (int)cudaGetLastError();
}

/* workspace: channels * H * W floats (needed by the kernel when channels > MAX_REGISTER_CHANNELS = 3) */
int ref_nd_rasterize_sum_backward(int n, int H, int W, int channels, const int *gaussian_ids_sorted,
                                  const int *tile_bins, const float *xys, const float *conics, const float *colors,
                                  const float *opacities, const float *background, const float *final_Ts,
                                  const int *final_idx, const float *v_output, const float *v_output_alpha,
                                  float *v_xy, float *v_conic, float *v_colors, float *v_opacity, float *workspace,
                                  cudaStream_t s) {
    dim3 tb((W + 15) / 16, (H + 15) / 16, 1), block(16, 16, 1), img(W, H, 1);
    cudaMemsetAsync(v_xy, 0, sizeof(float) * 2 * n, s); cudaMemsetAsync(v_conic, 0, sizeof(float) * 3 * n, s);
    cudaMemsetAsync(v_colors, 0, sizeof(float) * channels * n, s); cudaMemsetAsync(v_opacity, 0, sizeof(float) * n, s);
    cudaMemsetAsync(workspace, 0, sizeof(float) * channels * H * W, s);
    nd_rasterize_backward_sum_kernel<<<tb, block, 0, s>>>(
        tb, img, (unsigned)channels, gaussian_ids_sorted, (const int2 *)tile_bins, (const float2 *)xys,
        (const float3 *)conics, colors, opacities, background, final_Ts, final_idx, v_output, v_output_alpha,
        (float2 *)v_xy, (float3 *)v_conic, v_colors, v_opacity, workspace);
    return (int)cudaGetLastError();
}

} /* extern "C" */
