/*
 * gs2d_oracle.c -- CPU restatement of the gsplat 2D-Gaussian hot path of ac-freeman/GSVC.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (gsvc_b200/) links, imports or
 * calls this file.  Only tests/, __graft_entry__.smoke() and the cpu_baseline /
 * --impl reference legs of bench.py may use it, and only as the checker or the
 * timed CPU baseline -- never as the thing shipped.
 *
 * Every function restates one reference kernel / Python step and cites it
 * (paths relative to the reference tree, gsplat/gsplat/cuda/csrc/ unless noted).
 * Arithmetic is IEEE fp32 with contraction off (compile with -ffp-contract=off);
 * the reference GPU build uses FMA contraction, rcp/sqrt/ex2.approx (setup.py
 * --use_fast_math), so float outputs agree to tolerance (images 1e-5, gradients
 * 1e-4 relative) and integer outputs agree except at measure-zero rounding
 * boundaries (ceil / trunc / alpha-threshold flips), which the tests count.
 *
 * Parity pinning: compute_cov2d_bounds, get_tile_bbox, map_gaussian_to_intersects
 * and get_tile_bin_edges are pinned against golden vectors produced by the
 * reference's own Python restatement (gsplat/gsplat/_torch_impl.py) with the
 * inputs of the reference's tests (tests/golden/make_golden.py).  The 2D
 * projection, the sum rasteriser and all backwards have no golden vector, test or
 * CPU implementation anywhere in the reference ("parity unpinned" by reference
 * fixtures); they are pinned on the GPU box against the reference's own CUDA
 * kernels compiled from /root/reference into oracle/_ref/ (oracle/build_ref.sh).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GS_BLOCK 16        /* config.h:1-2  BLOCK_X = BLOCK_Y = 16 */
#define GS_BLOCK_SIZE 256  /* config.h:3    BLOCK_SIZE            */

/* ---- column-major 2x2 matrix with GLM semantics: m.c[col][row] ------------------ */
typedef struct { float c[2][2]; } mat2;

static mat2 m2(float a, float b, float c, float d) { /* glm::mat2(x0,y0,x1,y1) */
    mat2 m; m.c[0][0] = a; m.c[0][1] = b; m.c[1][0] = c; m.c[1][1] = d; return m;
}
static mat2 m2_mul(mat2 a, mat2 b) { /* glm operator*(mat2, mat2) */
    mat2 r;
    r.c[0][0] = a.c[0][0] * b.c[0][0] + a.c[1][0] * b.c[0][1];
    r.c[0][1] = a.c[0][1] * b.c[0][0] + a.c[1][1] * b.c[0][1];
    r.c[1][0] = a.c[0][0] * b.c[1][0] + a.c[1][0] * b.c[1][1];
    r.c[1][1] = a.c[0][1] * b.c[1][0] + a.c[1][1] * b.c[1][1];
    return r;
}
static mat2 m2_T(mat2 a) { return m2(a.c[0][0], a.c[1][0], a.c[0][1], a.c[1][1]); }
static mat2 m2_add(mat2 a, mat2 b) {
    return m2(a.c[0][0] + b.c[0][0], a.c[0][1] + b.c[0][1], a.c[1][0] + b.c[1][0], a.c[1][1] + b.c[1][1]);
}
static mat2 m2_neg(mat2 a) { return m2(-a.c[0][0], -a.c[0][1], -a.c[1][0], -a.c[1][1]); }

/* helpers.cuh:266-285 rotmat2d: glm::mat2(cos) then R[0][1] = -sin, R[1][0] = sin */
static mat2 rotmat2d(float rot) { float c = cosf(rot), s = sinf(rot); return m2(c, -s, s, c); }
/* helpers.cuh:287-298 rotmat2d_gradient */
static mat2 rotmat2d_gradient(float rot) { float c = cosf(rot), s = sinf(rot); return m2(-s, -c, c, -s); }
/* helpers.cuh:266-272 scale_to_mat2d */
static mat2 scale_to_mat2d(float sx, float sy) { return m2(sx, 0.f, 0.f, sy); }

/* helpers.cuh:45-68 compute_cov2d_bounds */
static int cov2d_bounds(const float cov[3], float conic[3], float *radius) {
    float det = cov[0] * cov[2] - cov[1] * cov[1];
    if (det == 0.f) return 0;
    float inv_det = 1.f / det;
    conic[0] = cov[2] * inv_det;
    conic[1] = -cov[1] * inv_det;
    conic[2] = cov[0] * inv_det;
    float b = 0.5f * (cov[0] + cov[2]);
    float disc = fmaxf(0.1f, b * b - det);
    float v1 = b + sqrtf(disc);
    float v2 = b - sqrtf(disc);
    *radius = ceilf(3.f * sqrtf(fmaxf(v1, v2)));
    return 1;
}

/* helpers.cuh:11-43 get_bbox / get_tile_bbox.  (int) truncates toward zero; the
 * clamp against the unsigned tile bound follows min(max(0,(int)v), bound). */
static int clampi(float v, int hi) {
    int i = (int)v;
    if (i < 0) i = 0;
    return i < hi ? i : hi;
}
static void tile_bbox(float cx, float cy, float radius, int tx, int ty, int mn[2], int mx[2]) {
    float tcx = cx / (float)GS_BLOCK, tcy = cy / (float)GS_BLOCK;
    float trx = radius / (float)GS_BLOCK, try_ = radius / (float)GS_BLOCK;
    mn[0] = clampi(tcx - trx, tx);
    mx[0] = clampi(tcx + trx + 1, tx);
    mn[1] = clampi(tcy - try_, ty);
    mx[1] = clampi(tcy + try_ + 1, ty);
}

/* bindings.cu:21-39 compute_cov2d_bounds_kernel (test utility).  Outputs are left
 * untouched (zero from the caller) when det == 0, as the device function returns
 * before writing and the kernel stores uninitialised registers; the reference test
 * masks those rows (tests/test_cov2d_bounds.py:33-34). */
void gs2d_oracle_cov2d_bounds(int n, const float *cov2d, float *conics, float *radii) {
    for (int i = 0; i < n; ++i) {
        float conic[3] = {0, 0, 0}, r = 0.f;
        cov2d_bounds(cov2d + 3 * i, conic, &r);
        conics[3 * i] = conic[0]; conics[3 * i + 1] = conic[1]; conics[3 * i + 2] = conic[2];
        radii[i] = r;
    }
}

/* get_tile_bbox exposed for the golden-vector test (helpers.cuh:27-43) */
void gs2d_oracle_tile_bbox(int n, const float *xys, const float *radii_f, int tx, int ty,
                           int *tile_min, int *tile_max) {
    for (int i = 0; i < n; ++i)
        tile_bbox(xys[2 * i], xys[2 * i + 1], radii_f[i], tx, ty, tile_min + 2 * i, tile_max + 2 * i);
}

/* shared tail of both projection forwards: foward2d.cu:49-67 / :104-122 */
static void project_tail(int i, float cx, float cy, const float cov[3], int tx, int ty,
                         float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit) {
    float conic[3], radius;
    if (!cov2d_bounds(cov, conic, &radius)) return; /* zero determinant */
    conics[3 * i] = conic[0]; conics[3 * i + 1] = conic[1]; conics[3 * i + 2] = conic[2];
    xys[2 * i] = cx; xys[2 * i + 1] = cy;
    radii[i] = (int)radius;
    int mn[2], mx[2];
    tile_bbox(cx, cy, radius, tx, ty, mn, mx);
    int area = (mx[0] - mn[0]) * (mx[1] - mn[1]);
    if (area <= 0) return;
    num_tiles_hit[i] = area;
    depths[i] = 0.0f;
}

/* foward2d.cu:12-69 project_gaussians_2d_forward_kernel.  All outputs start at 0
 * (bindings.cu:808-817 torch::zeros). */
void gs2d_oracle_project_chol_fwd(int n, const float *means2d, const float *L, int H, int W,
                                  int tx, int ty, float *xys, float *depths, int *radii,
                                  float *conics, int *num_tiles_hit) {
    memset(xys, 0, sizeof(float) * 2 * n); memset(depths, 0, sizeof(float) * n);
    memset(radii, 0, sizeof(int) * n); memset(conics, 0, sizeof(float) * 3 * n);
    memset(num_tiles_hit, 0, sizeof(int) * n);
    for (int i = 0; i < n; ++i) {
        float cx = 0.5f * (float)W * means2d[2 * i] + 0.5f * (float)W;
        float cy = 0.5f * (float)H * means2d[2 * i + 1] + 0.5f * (float)H;
        float l11 = L[3 * i], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
        float cov[3] = {l11 * l11, l11 * l21, l21 * l21 + l22 * l22};
        project_tail(i, cx, cy, cov, tx, ty, xys, depths, radii, conics, num_tiles_hit);
    }
}

/* foward2d.cu:71-124 project_gaussians_2d_scale_rot_forward_kernel */
void gs2d_oracle_project_scalerot_fwd(int n, const float *means2d, const float *scales,
                                      const float *rot, int H, int W, int tx, int ty, float *xys,
                                      float *depths, int *radii, float *conics, int *num_tiles_hit) {
    memset(xys, 0, sizeof(float) * 2 * n); memset(depths, 0, sizeof(float) * n);
    memset(radii, 0, sizeof(int) * n); memset(conics, 0, sizeof(float) * 3 * n);
    memset(num_tiles_hit, 0, sizeof(int) * n);
    for (int i = 0; i < n; ++i) {
        float cx = 0.5f * (float)W * means2d[2 * i] + 0.5f * (float)W;
        float cy = 0.5f * (float)H * means2d[2 * i + 1] + 0.5f * (float)H;
        mat2 R = rotmat2d(rot[i]);
        mat2 S = scale_to_mat2d(scales[2 * i], scales[2 * i + 1]);
        mat2 M = m2_mul(R, S);
        mat2 t = m2_mul(M, m2_T(M));
        float cov[3] = {t.c[0][0], t.c[0][1], t.c[1][1]};
        project_tail(i, cx, cy, cov, tx, ty, xys, depths, radii, conics, num_tiles_hit);
    }
}

/* gsplat/gsplat/utils.py:116-118 compute_cumulative_intersects: inclusive int32 cumsum */
int64_t gs2d_oracle_cumsum(int n, const int *num_tiles_hit, int *cum) {
    int32_t acc = 0;
    for (int i = 0; i < n; ++i) { acc += num_tiles_hit[i]; cum[i] = acc; }
    return n > 0 ? (int64_t)acc : 0;
}

/* forward.cu:100-136 map_gaussian_to_intersects.  The key is
 * (tile_id << 32) | (int64)(int32 bits of depth) -- note the sign extension. */
void gs2d_oracle_map_intersects(int n, const float *xys, const float *depths, const int *radii,
                                const int *cum, int tx, int ty, int64_t *isect_ids, int *gaussian_ids) {
    for (int i = 0; i < n; ++i) {
        if (radii[i] <= 0) continue;
        int mn[2], mx[2];
        tile_bbox(xys[2 * i], xys[2 * i + 1], (float)radii[i], tx, ty, mn, mx);
        int32_t cur = (i == 0) ? 0 : cum[i - 1];
        int32_t dbits; memcpy(&dbits, depths + i, 4);
        int64_t depth_id = (int64_t)dbits;
        for (int y = mn[1]; y < mx[1]; ++y)
            for (int x = mn[0]; x < mx[0]; ++x) {
                int64_t tile_id = (int64_t)(y * tx + x);
                isect_ids[cur] = (tile_id << 32) | depth_id;
                gaussian_ids[cur] = i;
                ++cur;
            }
    }
}

/* gsplat/gsplat/utils.py:164-165: torch.sort(int64) + gather.  Stable ascending
 * (tie order = ascending unsorted position).  LSD radix, 8 bits/pass,
 * signed order via flipping the sign bit. */
void gs2d_oracle_sort(int n, const int64_t *keys, const int *vals, int64_t *keys_out, int *vals_out) {
    if (n <= 0) return;
    uint64_t *ka = (uint64_t *)malloc(sizeof(uint64_t) * n), *kb = (uint64_t *)malloc(sizeof(uint64_t) * n);
    int *va = (int *)malloc(sizeof(int) * n), *vb = (int *)malloc(sizeof(int) * n);
    for (int i = 0; i < n; ++i) { ka[i] = (uint64_t)keys[i] ^ 0x8000000000000000ull; va[i] = vals[i]; }
    for (int pass = 0; pass < 8; ++pass) {
        size_t cnt[257]; memset(cnt, 0, sizeof(cnt));
        int sh = pass * 8;
        for (int i = 0; i < n; ++i) cnt[((ka[i] >> sh) & 255) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int i = 0; i < n; ++i) { size_t p = cnt[(ka[i] >> sh) & 255]++; kb[p] = ka[i]; vb[p] = va[i]; }
        uint64_t *tk = ka; ka = kb; kb = tk; int *tv = va; va = vb; vb = tv;
    }
    for (int i = 0; i < n; ++i) { keys_out[i] = (int64_t)(ka[i] ^ 0x8000000000000000ull); vals_out[i] = va[i]; }
    free(ka); free(kb); free(va); free(vb);
}

/* forward.cu:141-163 get_tile_bin_edges.  tile_bins has `rows` rows of (lo,hi),
 * zero-initialised by the caller (bindings.cu:319-321 allocates num_intersects
 * rows -- SURVEY Q6; rows must cover every tile id that occurs). */
void gs2d_oracle_tile_bin_edges(int n, const int64_t *keys_sorted, int *tile_bins) {
    for (int i = 0; i < n; ++i) {
        int32_t cur = (int32_t)(keys_sorted[i] >> 32);
        if (i == 0) tile_bins[2 * cur] = 0;
        if (i == n - 1) tile_bins[2 * cur + 1] = n;
        if (i == 0) continue;
        int32_t prev = (int32_t)(keys_sorted[i - 1] >> 32);
        if (prev != cur) { tile_bins[2 * prev + 1] = i; tile_bins[2 * cur] = i; }
    }
}

/* forward.cu:590-602: per pixel x Gaussian evaluation shared by forward and backward.
 * returns 1 if the pair passes (sigma >= 0 and alpha >= 1/255). */
static inline int eval_pair(const float *xy, const float *conic, float opac, float px, float py,
                            float *dx, float *dy, float *vis, float *alpha) {
    *dx = xy[0] - px; *dy = xy[1] - py;
    float sigma = 0.5f * (conic[0] * *dx * *dx + conic[2] * *dy * *dy) + conic[1] * *dx * *dy;
    *vis = expf(-sigma);
    *alpha = fminf(1.f, opac * *vis);
    if (sigma < 0.f || *alpha < 1.f / 255.f) return 0;
    return 1;
}

/* forward.cu:512-627 rasterize_forward_sum.  Only batch 0 of each tile is rendered:
 * `done = true` at :613 makes the loop leave after the first 256 entries.
 * final_Ts = 1, final_idx = sorted index of the last passing entry (initially 0),
 * no background. */
void gs2d_oracle_rasterize_sum_fwd(int H, int W, int tx, int ty, const int *ids_sorted,
                                   const int *tile_bins, const float *xys, const float *conics,
                                   const float *colors, const float *opac, float *out_img,
                                   float *final_Ts, int *final_idx) {
#pragma omp parallel for schedule(dynamic, 8)
    for (int t = 0; t < tx * ty; ++t) {
        int tile_y = t / tx, tile_x = t % tx;
        int lo = tile_bins[2 * t], hi = tile_bins[2 * t + 1];
        int nb = hi - lo; if (nb > GS_BLOCK_SIZE) nb = GS_BLOCK_SIZE; if (nb < 0) nb = 0;
        for (int ly = 0; ly < GS_BLOCK; ++ly) {
            int i = tile_y * GS_BLOCK + ly; if (i >= H) break;
            for (int lx = 0; lx < GS_BLOCK; ++lx) {
                int j = tile_x * GS_BLOCK + lx; if (j >= W) break;
                float px = (float)j, py = (float)i;
                float pix[3] = {0.f, 0.f, 0.f};
                int cur_idx = 0;
                for (int k = 0; k < nb; ++k) {
                    int g = ids_sorted[lo + k];
                    float dx, dy, vis, alpha;
                    if (!eval_pair(xys + 2 * g, conics + 3 * g, opac[g], px, py, &dx, &dy, &vis, &alpha)) continue;
                    pix[0] = pix[0] + colors[3 * g] * alpha;
                    pix[1] = pix[1] + colors[3 * g + 1] * alpha;
                    pix[2] = pix[2] + colors[3 * g + 2] * alpha;
                    cur_idx = lo + k;
                }
                size_t p = (size_t)i * W + j;
                final_Ts[p] = 1.f;
                final_idx[p] = cur_idx;
                out_img[3 * p] = pix[0]; out_img[3 * p + 1] = pix[1]; out_img[3 * p + 2] = pix[2];
            }
        }
    }
}

/* backward.cu:696-862 rasterize_backward_sum_kernel.  Every sorted entry of the tile
 * with index <= final_idx[pixel] that passes the alpha test receives gradient
 * (:784-805).  Accumulation here is in double per Gaussian (the reference sums fp32
 * partials with warp reductions + atomicAdd in nondeterministic order); the result
 * is rounded to fp32 once.  Outputs: v_xy[N,2], v_conic[N,3], v_rgb[N,3], v_opacity[N]. */
void gs2d_oracle_rasterize_sum_bwd(int n, int H, int W, int tx, int ty, const int *ids_sorted,
                                   const int *tile_bins, const float *xys, const float *conics,
                                   const float *colors, const float *opac, const int *final_idx,
                                   const float *v_out, float *v_xy, float *v_conic, float *v_rgb,
                                   float *v_opacity) {
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    double *acc = (double *)calloc((size_t)nthreads * n * 9, sizeof(double));
#pragma omp parallel
    {
        int tid = 0;
#ifdef _OPENMP
        tid = omp_get_thread_num();
#endif
        double *a = acc + (size_t)tid * n * 9;
#pragma omp for schedule(static)
        for (int t = 0; t < tx * ty; ++t) {
            int tile_y = t / tx, tile_x = t % tx;
            int lo = tile_bins[2 * t], hi = tile_bins[2 * t + 1];
            for (int ly = 0; ly < GS_BLOCK; ++ly) {
                int i = tile_y * GS_BLOCK + ly; if (i >= H) break;
                for (int lx = 0; lx < GS_BLOCK; ++lx) {
                    int j = tile_x * GS_BLOCK + lx; if (j >= W) break;
                    size_t p = (size_t)i * W + j;
                    float px = (float)j, py = (float)i;
                    int bin_final = final_idx[p];
                    const float *vo = v_out + 3 * p;
                    int end = hi - 1 < bin_final ? hi - 1 : bin_final;
                    for (int k = lo; k <= end; ++k) {
                        int g = ids_sorted[k];
                        float dx, dy, vis, alpha;
                        const float *c = conics + 3 * g;
                        if (!eval_pair(xys + 2 * g, c, opac[g], px, py, &dx, &dy, &vis, &alpha)) continue;
                        float v_alpha = 0.f;
                        v_alpha += colors[3 * g] * vo[0];
                        v_alpha += colors[3 * g + 1] * vo[1];
                        v_alpha += colors[3 * g + 2] * vo[2];
                        float v_sigma = -opac[g] * vis * v_alpha;
                        double *ag = a + (size_t)g * 9;
                        ag[0] += v_sigma * (c[0] * dx + c[1] * dy);      /* v_xy.x */
                        ag[1] += v_sigma * (c[1] * dx + c[2] * dy);      /* v_xy.y */
                        ag[2] += 0.5f * v_sigma * dx * dx;               /* v_conic.x */
                        ag[3] += 0.5f * v_sigma * dx * dy;               /* v_conic.y */
                        ag[4] += 0.5f * v_sigma * dy * dy;               /* v_conic.z */
                        ag[5] += alpha * vo[0];                          /* v_rgb */
                        ag[6] += alpha * vo[1];
                        ag[7] += alpha * vo[2];
                        ag[8] += vis * v_alpha;                          /* v_opacity */
                    }
                }
            }
        }
    }
    for (int g = 0; g < n; ++g) {
        double s[9] = {0};
        for (int th = 0; th < nthreads; ++th)
            for (int q = 0; q < 9; ++q) s[q] += acc[((size_t)th * n + g) * 9 + q];
        v_xy[2 * g] = (float)s[0]; v_xy[2 * g + 1] = (float)s[1];
        v_conic[3 * g] = (float)s[2]; v_conic[3 * g + 1] = (float)s[3]; v_conic[3 * g + 2] = (float)s[4];
        v_rgb[3 * g] = (float)s[5]; v_rgb[3 * g + 1] = (float)s[6]; v_rgb[3 * g + 2] = (float)s[7];
        v_opacity[g] = (float)s[8];
    }
    free(acc);
}

/* helpers.cuh:71-82 cov2d_to_conic_vjp */
static void cov2d_to_conic_vjp(const float conic[3], const float v_conic[3], float v_cov2d[3]) {
    mat2 X = m2(conic[0], conic[1], conic[1], conic[2]);
    mat2 G = m2(v_conic[0], v_conic[1], v_conic[1], v_conic[2]);
    mat2 vS = m2_mul(m2_mul(m2_neg(X), G), X);
    v_cov2d[0] = vS.c[0][0];
    v_cov2d[1] = vS.c[1][0] + vS.c[0][1];
    v_cov2d[2] = vS.c[1][1];
}

/* backward2d.cu:8-51 project_gaussians_2d_backward_kernel.  Reference formula, not
 * the analytic gradient: the off-diagonal term carries a factor 2.
 * Outputs start at zero and stay zero where radii <= 0. */
void gs2d_oracle_project_chol_bwd(int n, const float *L, int H, int W, const int *radii,
                                  const float *conics, const float *v_xy, const float *v_conic,
                                  float *v_cov2d, float *v_mean2d, float *v_L) {
    memset(v_cov2d, 0, sizeof(float) * 3 * n); memset(v_mean2d, 0, sizeof(float) * 2 * n);
    memset(v_L, 0, sizeof(float) * 3 * n);
    for (int i = 0; i < n; ++i) {
        if (radii[i] <= 0) continue;
        cov2d_to_conic_vjp(conics + 3 * i, v_conic + 3 * i, v_cov2d + 3 * i);
        float G11 = v_cov2d[3 * i], G12 = v_cov2d[3 * i + 1], G22 = v_cov2d[3 * i + 2];
        float l11 = L[3 * i], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
        v_L[3 * i] = 2 * l11 * G11 + 2 * G12 * l21;
        v_L[3 * i + 1] = 2 * l11 * G12 + 2 * l21 * G22;
        v_L[3 * i + 2] = 2 * l22 * G22;
        v_mean2d[2 * i] = v_xy[2 * i] * (0.5f * (float)W);
        v_mean2d[2 * i + 1] = v_xy[2 * i + 1] * (0.5f * (float)H);
    }
}

/* backward2d.cu:53-110 project_gaussians_2d_scale_rot_backward_kernel (factor 2 on
 * G_12 as in the reference). */
void gs2d_oracle_project_scalerot_bwd(int n, const float *scales, const float *rot, int H, int W,
                                      const int *radii, const float *conics, const float *v_xy,
                                      const float *v_conic, float *v_cov2d, float *v_mean2d,
                                      float *v_scale, float *v_rot) {
    memset(v_cov2d, 0, sizeof(float) * 3 * n); memset(v_mean2d, 0, sizeof(float) * 2 * n);
    memset(v_scale, 0, sizeof(float) * 2 * n); memset(v_rot, 0, sizeof(float) * n);
    for (int i = 0; i < n; ++i) {
        if (radii[i] <= 0) continue;
        cov2d_to_conic_vjp(conics + 3 * i, v_conic + 3 * i, v_cov2d + 3 * i);
        mat2 R = rotmat2d(rot[i]);
        mat2 Rg = rotmat2d_gradient(rot[i]);
        mat2 S = scale_to_mat2d(scales[2 * i], scales[2 * i + 1]);
        mat2 M = m2_mul(R, S);
        mat2 theta_g = m2_add(m2_mul(m2_mul(Rg, S), m2_T(M)), m2_mul(m2_mul(M, m2_T(S)), m2_T(Rg)));
        mat2 sxg = m2(2.f * scales[2 * i], 0.f, 0.f, 0.f);
        mat2 syg = m2(0.f, 0.f, 0.f, 2.f * scales[2 * i + 1]);
        mat2 sigx = m2_mul(m2_mul(R, sxg), m2_T(R));
        mat2 sigy = m2_mul(m2_mul(R, syg), m2_T(R));
        float G11 = v_cov2d[3 * i], G12 = v_cov2d[3 * i + 1], G22 = v_cov2d[3 * i + 2];
        v_scale[2 * i] = G11 * sigx.c[0][0] + 2 * G12 * sigx.c[0][1] + G22 * sigx.c[1][1];
        v_scale[2 * i + 1] = G11 * sigy.c[0][0] + 2 * G12 * sigy.c[0][1] + G22 * sigy.c[1][1];
        v_rot[i] = G11 * theta_g.c[0][0] + 2 * G12 * theta_g.c[0][1] + G22 * theta_g.c[1][1];
        v_mean2d[2 * i] = v_xy[2 * i] * (0.5f * (float)W);
        v_mean2d[2 * i + 1] = v_xy[2 * i + 1] * (0.5f * (float)H);
    }
}

int gs2d_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py's CPU arm: use this many OpenMP threads from now on (torchrun exports OMP_NUM_THREADS=1
 * to every rank, which would otherwise leave the timed CPU baseline on one core) */
void gs2d_oracle_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
