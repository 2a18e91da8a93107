// project.cu -- 2D projection of Cholesky / scale-rotation Gaussians, forward and backward.
// One thread per Gaussian; every output element is written (no pre-zeroing pass).
// Reference: foward2d.cu:12-124, backward2d.cu:8-110, helpers.cuh:11-82,266-298.
#include "gs2d_host.h"
#include "gs2d_math.cuh"

namespace gs2d {

constexpr int kProjThreads = 256;

// shared tail of both forwards (reference foward2d.cu:49-67 / :104-122)
__device__ __forceinline__ void project_tail(int i, float cx, float cy, float a, float b, float c, int tx, int ty,
                                             float* __restrict__ xys, float* __restrict__ depths,
                                             int32_t* __restrict__ radii, float* __restrict__ conics,
                                             int32_t* __restrict__ num_tiles_hit) {
    float3 conic = make_float3(0.f, 0.f, 0.f);
    float2 xy = make_float2(0.f, 0.f);
    float radius = 0.f;
    int r = 0, nth = 0;
    if (cov2d_bounds(a, b, c, conic, radius)) {
        xy = make_float2(cx, cy);
        r = trunc_ftz(radius);
        int x0, y0, x1, y1;
        tile_bbox(cx, cy, radius, tx, ty, x0, y0, x1, y1);
        const int area = (x1 - x0) * (y1 - y0);
        nth = area > 0 ? area : 0;
    } else {
        conic = make_float3(0.f, 0.f, 0.f);   // zero determinant: reference leaves the zero-init
    }
    reinterpret_cast<float2*>(xys)[i] = xy;
    conics[3 * i + 0] = conic.x; conics[3 * i + 1] = conic.y; conics[3 * i + 2] = conic.z;
    radii[i] = r;
    num_tiles_hit[i] = nth;
    depths[i] = 0.f;
}

__global__ void __launch_bounds__(kProjThreads)
k_project_chol_fwd(int n, const float* __restrict__ means2d, const float* __restrict__ L, float half_w, float half_h,
                   int tx, int ty, float* __restrict__ xys, float* __restrict__ depths, int32_t* __restrict__ radii,
                   float* __restrict__ conics, int32_t* __restrict__ num_tiles_hit) {
    const int i = blockIdx.x * kProjThreads + threadIdx.x;
    if (i >= n) return;
    const float2 m = reinterpret_cast<const float2*>(means2d)[i];
    const float cx = fmaf_(half_w, m.x, half_w);   // 0.5*W*x + 0.5*W  (foward2d.cu:37)
    const float cy = fmaf_(half_h, m.y, half_h);
    const float l11 = L[3 * i], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
    const float a = mulf(l11, l11);
    const float b = mulf(l11, l21);
    const float c = fmaf_(l21, l21, mulf(l22, l22));   // foward2d.cu:47
    project_tail(i, cx, cy, a, b, c, tx, ty, xys, depths, radii, conics, num_tiles_hit);
}

__global__ void __launch_bounds__(kProjThreads)
k_project_scalerot_fwd(int n, const float* __restrict__ means2d, const float* __restrict__ scales,
                       const float* __restrict__ rot, float half_w, float half_h, int tx, int ty,
                       float* __restrict__ xys, float* __restrict__ depths, int32_t* __restrict__ radii,
                       float* __restrict__ conics, int32_t* __restrict__ num_tiles_hit) {
    const int i = blockIdx.x * kProjThreads + threadIdx.x;
    if (i >= n) return;
    const float2 m = reinterpret_cast<const float2*>(means2d)[i];
    const float cx = fmaf_(half_w, m.x, half_w);
    const float cy = fmaf_(half_h, m.y, half_h);
    const float2 s = reinterpret_cast<const float2*>(scales)[i];
    const float th = rot[i];
    const float cr = cos_approx(th), sr = sin_approx(th);
    // M = R*S in GLM column-major order with R = [c -s; s c] (helpers.cuh:274-285), S = diag(s)
    // (foward2d.cu:93-97).  The reference keeps the 0*x terms (no finite-math assumption).
    const float zc = mulf(0.f, cr), zs = mulf(0.f, sr);
    const float m00 = fmaf_(cr, s.x, zs);
    const float m01 = fmaf_(-sr, s.x, zc);
    const float m10 = fmaf_(sr, s.y, zc);
    const float m11 = fmaf_(cr, s.y, -zs);
    const float a = fmaf_(m00, m00, mulf(m10, m10));
    const float b = fmaf_(m00, m01, mulf(m10, m11));
    const float c = fmaf_(m01, m01, mulf(m11, m11));
    project_tail(i, cx, cy, a, b, c, tx, ty, xys, depths, radii, conics, num_tiles_hit);
}

__global__ void __launch_bounds__(kProjThreads)
k_project_chol_bwd(int n, const float* __restrict__ L, float half_w, float half_h, const int32_t* __restrict__ radii,
                   const float* __restrict__ conics, const float* __restrict__ v_xy, int ld_xy,
                   const float* __restrict__ v_conic, int ld_conic, float* __restrict__ v_cov2d,
                   float* __restrict__ v_mean2d, float* __restrict__ v_L, float offdiag) {
    const int i = blockIdx.x * kProjThreads + threadIdx.x;
    if (i >= n) return;
    float3 vc = make_float3(0.f, 0.f, 0.f), vl = make_float3(0.f, 0.f, 0.f);
    float2 vm = make_float2(0.f, 0.f);
    if (radii[i] > 0) {   // backward2d.cu:23-25
        const float3 x = make_float3(conics[3 * i], conics[3 * i + 1], conics[3 * i + 2]);
        const float* gp = v_conic + (size_t)i * ld_conic;
        vc = conic_vjp(x, make_float3(gp[0], gp[1], gp[2]));
        const float l11 = L[3 * i], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
        // reference formula (backward2d.cu:39-41): offdiag = 2, off-diagonal counted twice
        vl = chol_vjp(l11, l21, l22, vc, offdiag);
        const float* xp = v_xy + (size_t)i * ld_xy;
        vm = make_float2(xp[0] * half_w, xp[1] * half_h);   // backward2d.cu:48-49
    }
    v_cov2d[3 * i] = vc.x; v_cov2d[3 * i + 1] = vc.y; v_cov2d[3 * i + 2] = vc.z;
    v_L[3 * i] = vl.x; v_L[3 * i + 1] = vl.y; v_L[3 * i + 2] = vl.z;
    reinterpret_cast<float2*>(v_mean2d)[i] = vm;
}

__global__ void __launch_bounds__(kProjThreads)
k_project_scalerot_bwd(int n, const float* __restrict__ scales, const float* __restrict__ rot, float half_w,
                       float half_h, const int32_t* __restrict__ radii, const float* __restrict__ conics,
                       const float* __restrict__ v_xy, int ld_xy, const float* __restrict__ v_conic, int ld_conic,
                       float* __restrict__ v_cov2d, float* __restrict__ v_mean2d, float* __restrict__ v_scale,
                       float* __restrict__ v_rot, float offdiag) {
    const int i = blockIdx.x * kProjThreads + threadIdx.x;
    if (i >= n) return;
    float3 vc = make_float3(0.f, 0.f, 0.f);
    float2 vm = make_float2(0.f, 0.f), vs = make_float2(0.f, 0.f);
    float vr = 0.f;
    if (radii[i] > 0) {
        const float3 x = make_float3(conics[3 * i], conics[3 * i + 1], conics[3 * i + 2]);
        const float* gp = v_conic + (size_t)i * ld_conic;
        vc = conic_vjp(x, make_float3(gp[0], gp[1], gp[2]));
        const float2 s = reinterpret_cast<const float2*>(scales)[i];
        const float th = rot[i];
        const float c = cos_approx(th), sn = sin_approx(th);
        // math (row, col) view: R = [[c, s], [-s, c]], R' = [[-s, c], [-c, -s]]  (helpers.cuh:274-298)
        // Sigma = R S^2 R^T.  dSigma/dsx = R diag(2 sx, 0) R^T, dSigma/dsy = R diag(0, 2 sy) R^T,
        // dSigma/dtheta = R' S (RS)^T + (RS) S^T R'^T                    (backward2d.cu:88-98)
        const float sx2 = 2.f * s.x, sy2 = 2.f * s.y;
        const float gx00 = sx2 * c * c, gx01 = -sx2 * c * sn, gx11 = sx2 * sn * sn;
        const float gy00 = sy2 * sn * sn, gy01 = sy2 * sn * c, gy11 = sy2 * c * c;
        const float sxx = s.x * s.x, syy = s.y * s.y;
        // A = R' S^2 R^T ; theta_g = A + A^T
        const float a00 = -sn * sxx * c + c * syy * sn;
        const float a01 = sn * sxx * sn + c * syy * c;
        const float a10 = -c * sxx * c - sn * syy * sn;
        const float a11 = c * sxx * sn - sn * syy * c;
        const float t00 = 2.f * a00, t01 = a01 + a10, t11 = 2.f * a11;
        // backward2d.cu:103-105: offdiag = 2 (factor 2 on G_12); 1 = analytic gradient
        vs.x = vc.x * gx00 + offdiag * vc.y * gx01 + vc.z * gx11;
        vs.y = vc.x * gy00 + offdiag * vc.y * gy01 + vc.z * gy11;
        vr = vc.x * t00 + offdiag * vc.y * t01 + vc.z * t11;
        const float* xp = v_xy + (size_t)i * ld_xy;
        vm = make_float2(xp[0] * half_w, xp[1] * half_h);
    }
    v_cov2d[3 * i] = vc.x; v_cov2d[3 * i + 1] = vc.y; v_cov2d[3 * i + 2] = vc.z;
    reinterpret_cast<float2*>(v_scale)[i] = vs;
    v_rot[i] = vr;
    reinterpret_cast<float2*>(v_mean2d)[i] = vm;
}

// reference bindings.cu:21-39 (test utility around helpers.cuh:45-68)
__global__ void __launch_bounds__(kProjThreads)
k_cov2d_bounds(int n, const float* __restrict__ cov2d, float* __restrict__ conics, float* __restrict__ radii) {
    const int i = blockIdx.x * kProjThreads + threadIdx.x;
    if (i >= n) return;
    float3 conic = make_float3(0.f, 0.f, 0.f);
    float radius = 0.f;
    cov2d_bounds(cov2d[3 * i], cov2d[3 * i + 1], cov2d[3 * i + 2], conic, radius);
    conics[3 * i] = conic.x; conics[3 * i + 1] = conic.y; conics[3 * i + 2] = conic.z;
    radii[i] = radius;
}

}  // namespace gs2d

using namespace gs2d;

extern "C" {

int gs2d_project_chol_fwd(int n, const float* means2d, const float* L, int img_h, int img_w, int tx, int ty,
                          float /*clip_thresh*/, float* xys, float* depths, int32_t* radii, float* conics,
                          int32_t* num_tiles_hit, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && tx > 0 && ty > 0, "project_chol_fwd: bad sizes");
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_PROJECT_FWD, (cudaStream_t)stream);
        k_project_chol_fwd<<<div_up(n, kProjThreads), kProjThreads, 0, (cudaStream_t)stream>>>(
        n, means2d, L, 0.5f * (float)(unsigned)img_w, 0.5f * (float)(unsigned)img_h, tx, ty, xys, depths, radii,
        conics, num_tiles_hit);
    }
    GS2D_CHECK_LAUNCH("project_chol_fwd");
    return GS2D_OK;
}

int gs2d_project_scalerot_fwd(int n, const float* means2d, const float* scales, const float* rot, int img_h,
                              int img_w, int tx, int ty, float /*clip_thresh*/, float* xys, float* depths,
                              int32_t* radii, float* conics, int32_t* num_tiles_hit, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && img_h > 0 && img_w > 0 && tx > 0 && ty > 0, "project_scalerot_fwd: bad sizes");
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_PROJECT_FWD, (cudaStream_t)stream);
        k_project_scalerot_fwd<<<div_up(n, kProjThreads), kProjThreads, 0, (cudaStream_t)stream>>>(
        n, means2d, scales, rot, 0.5f * (float)(unsigned)img_w, 0.5f * (float)(unsigned)img_h, tx, ty, xys, depths,
        radii, conics, num_tiles_hit);
    }
    GS2D_CHECK_LAUNCH("project_scalerot_fwd");
    return GS2D_OK;
}

static int project_chol_bwd(int n, const float* L, int img_h, int img_w, const int32_t* radii, const float* conics,
                            const float* v_xy, int ld_v_xy, const float* v_conic, int ld_v_conic, float* v_cov2d,
                            float* v_mean2d, float* v_L, float offdiag, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && ld_v_xy >= 2 && ld_v_conic >= 3, "project_chol_bwd: bad sizes");
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_PROJECT_BWD, (cudaStream_t)stream);
        k_project_chol_bwd<<<div_up(n, kProjThreads), kProjThreads, 0, (cudaStream_t)stream>>>(
        n, L, 0.5f * (float)(unsigned)img_w, 0.5f * (float)(unsigned)img_h, radii, conics, v_xy, ld_v_xy, v_conic,
        ld_v_conic, v_cov2d, v_mean2d, v_L, offdiag);
    }
    GS2D_CHECK_LAUNCH("project_chol_bwd");
    return GS2D_OK;
}

int gs2d_project_chol_bwd(int n, const float* L, int img_h, int img_w, const int32_t* radii, const float* conics,
                          const float* v_xy, int ld_v_xy, const float* v_conic, int ld_v_conic, float* v_cov2d,
                          float* v_mean2d, float* v_L, gs2d_stream_t stream) {
    return project_chol_bwd(n, L, img_h, img_w, radii, conics, v_xy, ld_v_xy, v_conic, ld_v_conic, v_cov2d, v_mean2d,
                            v_L, 2.f, stream);
}

int gs2d_project_chol_bwd_analytic(int n, const float* L, int img_h, int img_w, const int32_t* radii,
                                   const float* conics, const float* v_xy, int ld_v_xy, const float* v_conic,
                                   int ld_v_conic, float* v_cov2d, float* v_mean2d, float* v_L, gs2d_stream_t stream) {
    return project_chol_bwd(n, L, img_h, img_w, radii, conics, v_xy, ld_v_xy, v_conic, ld_v_conic, v_cov2d, v_mean2d,
                            v_L, 1.f, stream);
}

static int project_scalerot_bwd(int n, const float* scales, const float* rot, int img_h, int img_w,
                                const int32_t* radii, const float* conics, const float* v_xy, int ld_v_xy,
                                const float* v_conic, int ld_v_conic, float* v_cov2d, float* v_mean2d, float* v_scale,
                                float* v_rot, float offdiag, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && ld_v_xy >= 2 && ld_v_conic >= 3, "project_scalerot_bwd: bad sizes");
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_PROJECT_BWD, (cudaStream_t)stream);
        k_project_scalerot_bwd<<<div_up(n, kProjThreads), kProjThreads, 0, (cudaStream_t)stream>>>(
        n, scales, rot, 0.5f * (float)(unsigned)img_w, 0.5f * (float)(unsigned)img_h, radii, conics, v_xy, ld_v_xy,
        v_conic, ld_v_conic, v_cov2d, v_mean2d, v_scale, v_rot, offdiag);
    }
    GS2D_CHECK_LAUNCH("project_scalerot_bwd");
    return GS2D_OK;
}

int gs2d_project_scalerot_bwd(int n, const float* scales, const float* rot, int img_h, int img_w,
                              const int32_t* radii, const float* conics, const float* v_xy, int ld_v_xy,
                              const float* v_conic, int ld_v_conic, float* v_cov2d, float* v_mean2d, float* v_scale,
                              float* v_rot, gs2d_stream_t stream) {
    return project_scalerot_bwd(n, scales, rot, img_h, img_w, radii, conics, v_xy, ld_v_xy, v_conic, ld_v_conic,
                                v_cov2d, v_mean2d, v_scale, v_rot, 2.f, stream);
}

int gs2d_project_scalerot_bwd_analytic(int n, const float* scales, const float* rot, int img_h, int img_w,
                                       const int32_t* radii, const float* conics, const float* v_xy, int ld_v_xy,
                                       const float* v_conic, int ld_v_conic, float* v_cov2d, float* v_mean2d,
                                       float* v_scale, float* v_rot, gs2d_stream_t stream) {
    return project_scalerot_bwd(n, scales, rot, img_h, img_w, radii, conics, v_xy, ld_v_xy, v_conic, ld_v_conic,
                                v_cov2d, v_mean2d, v_scale, v_rot, 1.f, stream);
}

int gs2d_cov2d_bounds(int n, const float* cov2d, float* conics, float* radii, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0, "cov2d_bounds: bad size");
    if (n == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_OTHER, (cudaStream_t)stream);
        k_cov2d_bounds<<<div_up(n, kProjThreads), kProjThreads, 0, (cudaStream_t)stream>>>(n, cov2d, conics, radii);
    }
    GS2D_CHECK_LAUNCH("cov2d_bounds");
    return GS2D_OK;
}

}  // extern "C"
