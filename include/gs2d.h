/*
 * gs2d.h -- C ABI of libgs2d.so: the gsplat 2D-Gaussian hot path of ac-freeman/GSVC,
 * written from scratch for NVIDIA B200 (sm_100a).
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - `stream` is a cudaStream_t (NULL = legacy default stream); nothing synchronises
 *     the host and nothing allocates: outputs and workspaces are caller-owned;
 *   - the return value is GS2D_OK or an error code; gs2d_last_error() gives the text of
 *     the calling thread's last failure.  Launch errors are checked after every launch
 *     (the reference checks none);
 *   - layouts are the reference's: row-major contiguous fp32 / int32 AoS exactly as its
 *     torch tensors ([N,2] xys, [N,3] conics, ...).
 * Each function cites the reference interface it replaces (paths relative to
 * gsplat/gsplat/cuda/csrc/ of the reference tree unless noted).
 */
#ifndef GS2D_H
#define GS2D_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS2D_OK 0
#define GS2D_ERR_ARG 1
#define GS2D_ERR_CUDA 2

typedef void* gs2d_stream_t; /* cudaStream_t */
typedef void* gs2d_event_t;  /* cudaEvent_t  */

int gs2d_version(void);
const char* gs2d_last_error(void);
/* compute capability major*10+minor of `device`; the library only runs on 100 (sm_100a) */
int gs2d_device_arch(int device);

/* ---- instrumentation (no reference counterpart) ------------------------------------------
 * gs2d_launch_count: kernels this library has launched since it was loaded (all threads).
 * gs2d_prof_enable(1) makes every entry point bracket each of its stages with a CUDA event pair
 * on the launching stream; gs2d_prof_read synchronises those events, adds the elapsed time and
 * the number of spans per GS2D_STAGE_* into total_ms[] / spans[] (num_stages entries) and
 * forgets them.  Off by default; the timed `value` runs of bench.py keep it off. */
#define GS2D_STAGE_PROJECT_FWD 0
#define GS2D_STAGE_MEMSET 1
#define GS2D_STAGE_SCAN_MAP 2
#define GS2D_STAGE_SORT 3
#define GS2D_STAGE_EDGES_PACK 4
#define GS2D_STAGE_RASTER_FWD 5
#define GS2D_STAGE_RASTER_BWD 6
#define GS2D_STAGE_PROJECT_BWD 7
#define GS2D_STAGE_FIT_TILES 8
#define GS2D_STAGE_ADAN 9
#define GS2D_STAGE_OTHER 10
#define GS2D_NUM_STAGES 11
long long gs2d_launch_count(void);
int gs2d_prof_enable(int on);
int gs2d_prof_read(double* total_ms, int* spans, int num_stages);

/* ---- projection ------------------------------------------------------------------- */
/* replaces project_gaussians_2d_forward_tensor, bindings.cu:787-839 (kernel foward2d.cu:12-69).
 * Writes every element of all five outputs (no pre-zeroing needed). clip_thresh is accepted
 * and unused, as in the reference. */
int gs2d_project_chol_fwd(int n, const float* means2d, const float* L_elements, int img_h, int img_w,
                          int tiles_x, int tiles_y, float clip_thresh, float* xys, float* depths, int32_t* radii,
                          float* conics, int32_t* num_tiles_hit, gs2d_stream_t stream);
/* replaces project_gaussians_2d_scale_rot_forward_tensor, bindings.cu:841-899 (foward2d.cu:71-124) */
int gs2d_project_scalerot_fwd(int n, const float* means2d, const float* scales2d, const float* rotation, int img_h,
                              int img_w, int tiles_x, int tiles_y, float clip_thresh, float* xys, float* depths,
                              int32_t* radii, float* conics, int32_t* num_tiles_hit, gs2d_stream_t stream);
/* replaces project_gaussians_2d_backward_tensor, bindings.cu:902-949 (backward2d.cu:8-51).
 * v_xy / v_conic rows may be strided: ld_* = floats between consecutive rows (2 / 3 when
 * contiguous).  Reference formula incl. the doubled off-diagonal term. */
int gs2d_project_chol_bwd(int n, const float* L_elements, int img_h, int img_w, const int32_t* radii,
                          const float* conics, const float* v_xy, int ld_v_xy, const float* v_conic, int ld_v_conic,
                          float* v_cov2d, float* v_mean2d, float* v_L_elements, gs2d_stream_t stream);
/* replaces project_gaussians_2d_scale_rot_backward_tensor, bindings.cu:951-1004 (backward2d.cu:53-110) */
int gs2d_project_scalerot_bwd(int n, const float* scales2d, const float* rotation, int img_h, int img_w,
                              const int32_t* radii, const float* conics, const float* v_xy, int ld_v_xy,
                              const float* v_conic, int ld_v_conic, float* v_cov2d, float* v_mean2d, float* v_scale,
                              float* v_rot, gs2d_stream_t stream);
/* SURVEY 8f.4 option, no reference counterpart: the same two backwards with the ANALYTIC gradient of
 * the forward.  The reference counts the off-diagonal covariance term twice (v_cov2d.y is already
 * the total derivative w.r.t. the shared off-diagonal, helpers.cuh:80, and backward2d.cu:39-40 /
 * :103-105 multiply it by 2 again); these use the factor 1, which is what autograd of
 * the forward chain gives (tests/test_gpu_extensions.py checks them against fp64 autograd).
 * Everything else -- arguments, outputs, radii <= 0 handling -- is identical. */
int gs2d_project_chol_bwd_analytic(int n, const float* L_elements, int img_h, int img_w, const int32_t* radii,
                                   const float* conics, const float* v_xy, int ld_v_xy, const float* v_conic,
                                   int ld_v_conic, float* v_cov2d, float* v_mean2d, float* v_L_elements,
                                   gs2d_stream_t stream);
int gs2d_project_scalerot_bwd_analytic(int n, const float* scales2d, const float* rotation, int img_h, int img_w,
                                       const int32_t* radii, const float* conics, const float* v_xy, int ld_v_xy,
                                       const float* v_conic, int ld_v_conic, float* v_cov2d, float* v_mean2d,
                                       float* v_scale, float* v_rot, gs2d_stream_t stream);
/* replaces compute_cov2d_bounds_tensor, bindings.cu:41-61 (helpers.cuh:45-68); radii is fp32 [n] */
int gs2d_cov2d_bounds(int n, const float* cov2d, float* conics, float* radii, gs2d_stream_t stream);

/* ---- binning, step by step (the reference's public utilities) ----------------------- */
/* replaces torch.cumsum(num_tiles_hit, dtype=int32), gsplat/gsplat/utils.py:116.  Inclusive scan.
 * scratch: gs2d_scan_scratch_bytes(n) bytes.  The total (= cum[n-1]) is also written to
 * total_out (device, may be NULL) and total_host (pinned host memory, may be NULL). */
size_t gs2d_scan_scratch_bytes(int n);
int gs2d_cumsum_tiles_hit(int n, const int32_t* num_tiles_hit, int32_t* cum_tiles_hit, int32_t* total_out,
                          int32_t* total_host, void* scratch, size_t scratch_bytes, gs2d_stream_t stream);
/* gs2d_cumsum_tiles_hit WITHOUT caller scratch and without any allocation.  No counterpart in the reference:
 * added for callers that cannot allocate (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  The scan's look-back state lives in a file-scope
 * `static __device__` array of the library (128 KB, held by every GPU that loads the library until exit),
 * which bounds the size: n <= GS2D_CUMSUM_STATIC_MAX_N, larger n is rejected with GS2D_ERR_ARG and a
 * clear message.  The entry point calls no cudaMalloc / cudaFree / allocating library routine; it clears
 * the part of the array it uses (stream-ordered memset) and launches the same kernel as
 * gs2d_cumsum_tiles_hit, so the results are identical.  ALL calls on a device share the one array:
 * calls that may overlap on the GPU (different streams without ordering between them) race -- serialise
 * them, or use gs2d_cumsum_tiles_hit with scratch of your own for the concurrent ones. */
#define GS2D_CUMSUM_STATIC_MAX_N (1 << 22)
int gs2d_cumsum_tiles_hit_static(int n, const int32_t* num_tiles_hit, int32_t* cum_tiles_hit, int32_t* total_out,
                                 int32_t* total_host, gs2d_stream_t stream);
/* replaces map_gaussian_to_intersects_tensor, bindings.cu:274-313 (forward.cu:100-136).
 * Key = (tile_id << 32) | sign-extended depth bits.  Slots of Gaussians with radii <= 0 are
 * written as zero, matching the reference's zero-initialised outputs. */
int gs2d_map_intersects(int n, int num_intersects, const float* xys, const float* depths, const int32_t* radii,
                        const int32_t* cum_tiles_hit, int tiles_x, int tiles_y, int64_t* isect_ids,
                        int32_t* gaussian_ids, gs2d_stream_t stream);
/* replaces torch.sort(int64) + torch.gather, gsplat/gsplat/utils.py:164-165: stable ascending
 * onesweep LSD radix sort of (int64 key, int32 value) pairs over key bits [begin_bit, end_bit)
 * (signed order when end_bit == 64).  scratch: gs2d_sort_scratch_bytes(n). */
size_t gs2d_sort_scratch_bytes(int n);
int gs2d_sort_pairs(int n, const int64_t* keys_in, const int32_t* vals_in, int64_t* keys_out, int32_t* vals_out,
                    int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, gs2d_stream_t stream);
/* replaces get_tile_bin_edges_tensor, bindings.cu:315-330 (forward.cu:141-163).  tile_bins has
 * `rows` rows of (lo, hi) and is zero-filled here; rows must exceed every tile id. */
int gs2d_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted, int32_t* tile_bins, int rows,
                        gs2d_stream_t stream);

/* ---- rasteriser ---------------------------------------------------------------------- */
/* Layout of the caller-owned state buffer of one forward (kept until its backward). */
typedef struct {
    size_t total_bytes;
    size_t off_status;      /* int32[8]: [0] num_intersects, [1] overflow (num_intersects > cap),
                               [2] OR of all depth bits, [3] min(num_intersects, cap) */
    size_t off_tile_bins;   /* int32[tiles][2] */
    size_t off_ids_sorted;  /* int32[cap]  gaussian_ids_sorted */
    size_t off_keys_sorted; /* uint32[cap] tile id of each sorted intersection */
    size_t off_packed;      /* float4[cap][3] sorted packed copy of the Gaussian parameters */
    size_t off_scratch;     /* sort ping-pong, histograms, look-back state (dead after forward) */
    size_t zero_begin;      /* [zero_begin, zero_begin + zero_bytes) is memset by the library */
    size_t zero_bytes;
    int key_bits;           /* ceil(log2(tiles)) : width of the radix sort */
    int sort_passes;
} gs2d_state_layout;
int gs2d_state_layout_for(int n, int cap, int tiles_x, int tiles_y, gs2d_state_layout* out);

/* Fused, sync-free replacement of the whole body of _RasterizeGaussiansSum.forward
 * (gsplat/gsplat/rasterize_sum.py:93-186): compute_cumulative_intersects (utils.py:99-118),
 * bin_and_sort_gaussians (utils.py:121-167: map forward.cu:100-136, torch.sort, gather, bin
 * edges forward.cu:141-163) and rasterize_forward_sum (forward.cu:512-627, host bindings.cu:400-469).
 * depths may be NULL (= all zero, which is what both projections produce); otherwise the OR of
 * the depth bits is reported in status[2] and a non-zero value means the caller must use the
 * step-by-step path (64-bit keys).  cap = capacity in intersections of the state buffer; when
 * num_intersects > cap nothing past cap is written, status[1] is set and the caller retries with
 * a larger buffer.  status_host (pinned, int32[4], may be NULL) receives status[0..3] as soon as
 * the scan stage is done; scan_done (may be NULL) is recorded right after that stage so the host
 * can wait for exactly that and no more.  out_img [H,W,3] fp32; final_idx [H,W] int32 or NULL
 * (render-only).  final_Ts is the constant 1 in the sum variant and is not
 * materialised. */
int gs2d_rasterize_sum_fwd(int n, const float* xys, const float* depths, const int32_t* radii, const float* conics,
                           const int32_t* num_tiles_hit, const float* colors, const float* opacities, int img_h,
                           int img_w, int cap, void* state, size_t state_bytes, float* out_img, int32_t* final_idx,
                           int32_t* status_host, gs2d_event_t scan_done, gs2d_stream_t stream);

/* The rasteriser kernel alone on caller-provided binning (replaces rasterize_forward_sum_tensor,
 * bindings.cu:400-469).  tile_bins must have a row for every tile.  final_Ts [H,W] may be NULL. */
int gs2d_rasterize_sum_fwd_binned(int n, int num_intersects, const int32_t* gaussian_ids_sorted,
                                  const int32_t* tile_bins, const float* xys, const float* conics,
                                  const float* colors, const float* opacities, int img_h, int img_w, float* out_img,
                                  float* final_Ts, int32_t* final_idx, void* scratch, size_t scratch_bytes,
                                  gs2d_stream_t stream);
size_t gs2d_raster_binned_scratch_bytes(int num_intersects);

/* Replaces rasterize_backward_sum_tensor, bindings.cu:706-779 (kernel backward.cu:696-862), on
 * the state of the matching forward.  v_out [H,W,3] fp32 (contiguous).  grads: fp32 [n][12] rows
 * = v_xy(2) v_conic(3) v_colors(3) v_opacity(1) pad(3); zero-filled here, then one 3 x 16-byte
 * reduction per (Gaussian, tile) pair. */
int gs2d_rasterize_sum_bwd(int n, int img_h, int img_w, int cap, const void* state, size_t state_bytes,
                           const int32_t* final_idx, const float* v_out, float* grads, gs2d_stream_t stream);
/* Same on caller-provided binning (ids sorted + tile_bins), for parity tests of the kernel alone. */
int gs2d_rasterize_sum_bwd_binned(int n, int num_intersects, const int32_t* gaussian_ids_sorted,
                                  const int32_t* tile_bins, const float* xys, const float* conics,
                                  const float* colors, const float* opacities, int img_h, int img_w,
                                  const int32_t* final_idx, const float* v_out, float* grads, void* scratch,
                                  size_t scratch_bytes, gs2d_stream_t stream);

/* Sync-free, CUDA-graph-capturable forward / backward of rasterize_gaussians_sum on count / scatter
 * binning (the default of the Python operator; same results as gs2d_rasterize_sum_fwd / _bwd).
 * Replaces the same reference code (rasterize_sum.py:93-254, utils.py:99-167, forward.cu:100-163,
 * 512-627, backward.cu:696-862) but orders a tile's Gaussians inside the tile kernels instead of
 * sorting (tile, id) keys: per-tile counts by atomics + a last-block scan give the reference's
 * tile_bins, a scatter fills each range in arbitrary order, and the forward / backward kernels put
 * a tile's batch into ascending Gaussian-id order -- the order of the reference's torch.sort on
 * tile|depth keys when depths are zero, which is what both projections emit -- while
 * staging it into shared memory.  Three kernels (count+scan, scatter-pack, tiles) chained with
 * programmatic dependent launch instead of five, no key / id arrays, and nothing the host has to
 * wait for: the intersection count stays on the device.
 *   state       gs2d_opstate_bytes(cap, H, W) bytes, 256-byte aligned, kept for the backward;
 *               [0, zero) is cleared here (one memset);
 *   background  fp32[3] or NULL: with no intersection at all the image is the background colour,
 *               as rasterize_sum.py:121-129 returns it (decided on the device);
 *   mail_host   PINNED int32[8] or NULL: {num_intersects, overflow (num_intersects > cap),
 *               max per tile, 0, mail_seq} written by the scan's last block, mail_seq last.  The
 *               host reads it whenever it next cares (capacity hint, lazy overflow check);
 *   overflow    when num_intersects > cap nothing is scattered or rendered (out_img is left
 *               untouched) and the backward writes zero gradients: the caller must regrow and redo.
 * Depths are taken to be zero (callers with non-zero depths use the step-by-step path).  The
 * N-channel / untruncated options stay on gs2d_rasterize_sum_nd_*. */
size_t gs2d_opstate_bytes(int cap, int img_h, int img_w);
/* byte offsets of tile_off (int32[tiles+1]), the packed records and the status words inside it */
int gs2d_opstate_ranges(int cap, int img_h, int img_w, size_t* off_tile_off, size_t* off_packed, size_t* off_status);
int gs2d_rasterize_sum_fwd_ranked(int n, const float* xys, const int32_t* radii, const float* conics,
                                  const float* colors, const float* opacities, const float* background, int img_h,
                                  int img_w, int cap, void* state, size_t state_bytes, float* out_img,
                                  int32_t* final_idx, int32_t* mail_host, int mail_seq, gs2d_stream_t stream);
int gs2d_rasterize_sum_bwd_ranked(int n, int img_h, int img_w, int cap, const void* state, size_t state_bytes,
                                  const int32_t* final_idx, const float* v_out, float* grads, gs2d_stream_t stream);

/* N-channel and / or untruncated accumulated-sum rasteriser.  The reference's Python
 * dispatches colours with channels != 3 to nd_rasterize_sum_forward / _backward (rasterize_sum.py:
 * 147-150, 218-221), which its extension never binds although the kernels exist
 * (nd_rasterize_forward_sum forward.cu:630-710, nd_rasterize_backward_sum_kernel backward.cu:864-978);
 * and its 3-channel kernel renders only the first 256 Gaussians of a tile.  These two
 * entry points close both gaps; arguments as gs2d_rasterize_sum_fwd / _bwd plus
 *   channels      1..32; colors is [n][channels], out_img / v_out are [H][W][channels]
 *   max_per_tile  256: the reference's 3-channel truncation; 0: every Gaussian of a tile is rendered
 *                 (what the reference's N-channel kernel does), in batches of 256
 *   v_colors      [n][channels] colour gradients (zero-filled here); `grads` rows keep v_xy(2)
 *                 v_conic(3) . . . v_opacity(1), their colour slots stay zero.
 * Same pixel test and accumulation order as the 3-channel kernels: channels = 3 with max_per_tile =
 * 256 is bit-identical to gs2d_rasterize_sum_fwd.  final_idx keeps the 3-channel meaning.
 * The backward gives every rendered Gaussian its gradient (the reference's unlaunched N-channel
 * backward starts one entry below final_index and so skips each tile's last Gaussian). */
int gs2d_rasterize_sum_nd_fwd(int n, int channels, const float* xys, const float* depths, const int32_t* radii,
                              const float* conics, const int32_t* num_tiles_hit, const float* colors,
                              const float* opacities, int img_h, int img_w, int cap, int max_per_tile, void* state,
                              size_t state_bytes, float* out_img, int32_t* final_idx, int32_t* status_host,
                              gs2d_event_t scan_done, gs2d_stream_t stream);
int gs2d_rasterize_sum_nd_bwd(int n, int channels, int img_h, int img_w, int cap, int max_per_tile, const void* state,
                              size_t state_bytes, const float* colors, const int32_t* final_idx, const float* v_out,
                              float* grads, float* v_colors, gs2d_stream_t stream);

/* ---- fused fit iteration and render (callers either side of the path) ----
 * One frame model = raw parameters exactly as the reference's GaussianVideo_frame holds them
 * (GaussianSplats_Represent.py:28-45): xyz_raw [n,2] (means2d = tanh), chol_raw [n,3]
 * (L = chol_raw + [0.5, 0, 0.5]), features [n,3] (colours; rgb_W = 1, opacity = 1).
 * gs2d_fit_step runs ONE iteration of GaussianVideo_frame.train_iter
 * (GaussianSplats_Represent.py:191-207: forward :83-90, L2 loss utils.py:21-26, backward,
 * Adan step optimizer.py:125-236 with the reference's defaults) as three or four kernels on `stream`
 * (two with gs2d_fit_options.slots, gs2d_fit_step_band) with no host synchronisation, no allocation
 * and no memset: project+count+scan, scatter-pack, tile kernel (rasterise forward + clamp/L2 +
 * rasterise backward) and projection-backward+Adan.  Everything lives in caller-owned memory:
 *   work        gs2d_fit_layout_for(...).total_bytes, 256-byte aligned; gs2d_fit_init zeroes its
 *               control region once (and the Adan state);
 *   adan_state  fp32 [4][n][8]: m, n, d, previous gradient of (xyz 2, chol 3, features 3);
 *   target      the frame to fit: GS2D_TARGET_F32_CHW fp32 planar [3][H][W] in [0,1], or
 *               GS2D_TARGET_U8_HWC uint8 interleaved [H][W][3] (compared as value / 255), or
 *               GS2D_TARGET_I420 the raw frame as the reference's .yuv file holds it (planar
 *               YUV 4:2:0, 8 bit: H*W luma bytes, then (H/2)*(W/2) U, then V; H and W even),
 *               converted per pixel inside the tile kernel exactly as the reference's loader
 *               does (cv2.COLOR_YUV2RGB_I420, utils.py:134-156; see gs2d_i420_to_rgb) and
 *               compared as value / 255 -- half the bytes of the RGB frame to move and read.
 * precounted: every step ends by projecting the UPDATED parameters and counting/scanning their
 * tiles, i.e. it pre-computes the next step's binning; pass precounted = 1 when the previous call
 * on this workspace was gs2d_fit_step and neither the parameters nor the workspace were touched
 * since (a steady-state iteration is then three kernels), 0 otherwise (first step, after loading
 * parameters, after gs2d_render_chol on the same workspace): the count kernel then runs first.
 * step is 1-based (bias correction).  The step's loss SUM (sum of squared errors over 3*H*W
 * values, a double) accumulates in work + off_loss + 8 * (step % GS2D_FIT_LOSS_SLOTS); mean =
 * sum / (3*H*W).  status (int32[16] at off_status): [0] num_intersects of the last step,
 * [1] != 0: the step index at which num_intersects first exceeded cap -- from then on every step
 * is a no-op (parameters untouched) until the caller re-initialises a larger workspace,
 * [3] largest per-tile count of the last step.
 * Optional outputs (may be NULL): out_img fp32 planar [3][H][W], clamped to [0,1] (the
 * reference's render); param_grads_out fp32 [n][8] the gradients handed to Adan;
 * ids_sorted_out int32[cap] the Gaussian ids of every tile range in rendering order (-1 beyond
 * the first 256 of a tile). */
#define GS2D_FIT_LOSS_SLOTS 64
#define GS2D_TARGET_F32_CHW 0
#define GS2D_TARGET_U8_HWC 1
#define GS2D_TARGET_I420 2
#define GS2D_TARGET_TILED 3 /* what gs2d_fit_target_prepare wrote */
typedef struct {
    size_t total_bytes;
    size_t off_status;      /* int32[16] */
    size_t off_loss;        /* double[GS2D_FIT_LOSS_SLOTS] */
    size_t off_tile_count;  /* int32[tiles]   per-tile fill counters (zero between kernels) */
    size_t off_tile_off;    /* int32[tiles+1] exclusive scan of the per-tile counts = tile ranges */
    size_t off_grads;       /* fp32[n][12]    v_xy(2) v_conic(3) v_colors(3) v_opacity(1) pad(3) */
    size_t off_packed;      /* float4[cap][3] packed records, grouped by tile */
    size_t zero_bytes;      /* [0, zero_bytes) is what gs2d_fit_init clears */
    int tiles;
} gs2d_fit_layout;
/* What a step reports, written by its last kernel straight into PINNED HOST memory when
 * report_host != NULL (device-visible, e.g. cudaHostAlloc / torch pin_memory): the host learns
 * the loss without issuing a copy -- it only has to wait for the step (event or stream sync) or
 * poll `step`, which is written last. */
typedef struct {
    double loss_sum;         /* sum of squared errors over 3*H*W values */
    int32_t num_intersects;  /* of this step */
    int32_t overflow_step;   /* see status[1] */
    int32_t max_per_tile;
    int32_t step;            /* the step index this record belongs to */
} gs2d_fit_report;
int gs2d_fit_layout_for(int n, int cap, int img_h, int img_w, gs2d_fit_layout* out);
/* A frame is fitted for thousands of iterations (train_video_Represent.py:85-86): convert it ONCE into the
 * order the tile kernel reads it in -- fp32, each thread's pixels' (r, g, b) contiguous, tile after tile;
 * gs2d_fit_target_bytes(H, W) bytes, 16-byte aligned -- and pass that with GS2D_TARGET_TILED.  Same
 * arithmetic as the in-kernel decode of the other three formats (so the same loss, bit for bit), but the
 * iteration then spends no instructions on it and loads the frame with coalesced 16-byte accesses.
 * A stream of frames that are used once each is better off with the raw formats. */
size_t gs2d_fit_target_bytes(int img_h, int img_w);
int gs2d_fit_target_prepare(int img_h, int img_w, const void* target, int target_format, float* prepared,
                            gs2d_stream_t stream);
int gs2d_fit_init(int n, int cap, int img_h, int img_w, void* work, size_t work_bytes, float* adan_state,
                  gs2d_stream_t stream);
int gs2d_fit_step(int n, int cap, int img_h, int img_w, float* xyz_raw, float* chol_raw, float* features,
                  float* adan_state, const void* target, int target_format, void* work, size_t work_bytes, int step,
                  float lr, float beta1, float beta2, float beta3, float eps, float weight_decay, int precounted,
                  float* out_img, float* param_grads_out, int32_t* ids_sorted_out, gs2d_fit_report* report_host,
                  gs2d_stream_t stream);
/* The same step for a frame split across GPUs by tile rows (BASELINE config 5): this
 * rank bins and rasterises only tile rows [tile_row_begin, tile_row_end) -- projection, the
 * per-Gaussian update and the optimizer stay replicated -- and the step is issued in two phases
 * so the caller can sum the ranks' gradients in between:
 *     gs2d_fit_step_band(..., GS2D_FIT_PHASE_GRADS, ...);
 *     ncclAllReduce(work + off_grads, n * 12 floats, sum)      -- same stream
 *     gs2d_fit_step_band(..., GS2D_FIT_PHASE_UPDATE, ...);
 * The loss is normalised by the WHOLE frame (3*H*W) on every rank; in this mode each tile's loss is
 * also added into float 9 of the gradient buffer (pad of row 0), so the one all-reduce sums the
 * ranks' partial losses too and the report's loss_sum is the global one.  All ranks then hold
 * bit-identical parameters.  gs2d_fit_step == full band, both phases. */
#define GS2D_FIT_PHASE_GRADS 1
#define GS2D_FIT_PHASE_UPDATE 2
/* Optional: exchange the gradients through peer-mapped memory (NVLink / NVSwitch) instead of a
 * collective call.  The tile kernel accumulates into `accumulate` (this rank's buffer for this
 * step, fp32 [n][12], zero on entry); after a cross-rank barrier issued by the caller on the same
 * stream, the UPDATE phase reads row i of every buffer in `peers` (a DEVICE array of num_peers
 * device pointers, own buffer included, all mapped into this rank's address space) in list order
 * and sums them -- the all-reduce is that read, fused into the per-Gaussian update kernel -- and
 * re-zeroes row i of `zero`.  With two buffers per rank used alternately (accumulate = buf[step&1],
 * zero = buf[(step+1)&1]) one barrier per step suffices: a buffer is only re-zeroed two steps
 * after peers read it.  exchange == NULL: gradients live in the workspace (off_grads). */
typedef struct {
    float* accumulate;
    const float* const* peers;
    int num_peers;
    float* zero;
    const float* multicast;   /* != NULL: the NVLink-switch multicast address of this step's buffers
                               * (all ranks' `accumulate` bound to one multicast object): the update
                               * kernel sums the rows with multimem.ld_reduce -- the switch adds the
                               * ranks' copies and returns one row, so a rank reads n*48 bytes
                               * whatever the world size.  peers / num_peers are then ignored. */
    /* ---- sharded update (shard_end > shard_begin; needs tile slots and `peers`) ------------------
     * The replicated update does not scale: every rank reads every rank's rows and redoes the
     * per-Gaussian work of the whole model.  Here rank r owns Gaussians [shard_begin, shard_end):
     * its UPDATE phase reduces only those rows (a reduce-scatter fused into the kernel), runs Adan
     * on them (the optimizer state of a Gaussian lives on its owner only), and stores the new
     * parameters into EVERY rank's parameter tensors (an all-gather fused into the kernel: plain
     * NVLink stores, 32 bytes per Gaussian and peer).  After a second barrier the next GRADS phase
     * scatters the records of its own tile-row band from its now complete local copy.
     * Two cross-rank barriers per step (after GRADS, after UPDATE), issued by the caller.
     *   xyz_peers / chol_peers / feat_peers   DEVICE arrays of num_peers pointers: rank q's
     *                    xyz_raw / chol_raw / features tensors (this rank's own included)
     *   status_peers     DEVICE array of num_peers pointers: rank q's status words (its work +
     *                    layout.off_status): a tile that outgrows its slots is flagged on every
     *                    rank, so that all of them stop at the same step */
    int shard_begin, shard_end;
    float* const* xyz_peers;
    float* const* chol_peers;
    float* const* feat_peers;
    int32_t* const* status_peers;
    /* ---- remote scatter (optional, with a sharded update) -----------------------------------------
     * Without it every rank re-projects ALL Gaussians each step to find the ones that touch its band (the
     * part of a step that does not shrink with the number of ranks).  With it the update kernel, which
     * has its shard's new parameters in registers, writes the next step's packed records straight into
     * the slots of the rank that rasterises each tile: the exchange of the records is an all-to-all fused
     * into the update kernel (system-scope atomics for the slot claims, plain NVLink stores for the 48-byte
     * records), and a steady-state step has no scatter kernel at all.
     *   work_peers   DEVICE array of num_peers pointers: rank q's workspace (all of one layout)
     *   band_edges   DEVICE int32[num_peers + 1]: rank q rasterises tile rows [band_edges[q], band_edges[q+1])
     * The report's num_intersects then counts the pairs of this rank's SHARD (over all bands). */
    void* const* work_peers;
    const int32_t* band_edges;
    /* ---- cross-rank barrier issued by the library (optional; any peer exchange) -------------------
     * With `barrier_pads` set and both phases in one call, the barrier between GRADS and UPDATE (and,
     * for a sharded update, the one after UPDATE) is a one-warp kernel of this library on the same
     * stream: the whole step is then ONE host call, which matters once a step is down to ~100 us.
     *   barrier_pads   DEVICE array of num_peers pointers: rank q's pad, uint32[num_peers] in peer-mapped
     *                  memory, zero before the first step.  Barrier number e: rank r stores e into
     *                  pad_q[r] of every q (release, system scope), then waits until its own pad holds
     *                  >= e in every slot (acquire).  Epochs only grow, so nothing is ever reset.
     *   rank           this rank's index into the peer arrays
     *   barrier_epoch  number of the first barrier of this call (the caller counts: 1, 2, ... -- every
     *                  rank must issue the same sequence of calls); a sharded step uses two
     * A rank that waits longer than 10 s traps (a peer died): an error, not a hang. */
    uint32_t* const* barrier_pads;
    int rank;
    uint32_t barrier_epoch;
    /* ---- pushed reduce-scatter (optional, with a sharded update; needs gs2d_fit_options.params8; the newest
     * field, hence last: the layout before it is as it was) ----------
     * push_per > 0: Gaussian i is owned by rank i / push_per (so shard_begin / shard_end must be that rank's
     * [r * push_per, min((r + 1) * push_per, n))), and the TILE kernel adds every pair's gradient straight into
     * row i of the owner's buffer `peers[i / push_per]` -- reductions without a return value, fire and forget
     * over NVLink, 32 bytes per (Gaussian, tile) pair as two 16-byte vector reductions
     * (red.global.add.v4.f32).  After the barrier the update kernel finds its shard's rows complete in its OWN
     * buffer (`accumulate`): no peer reads (they made the sharded update's kernel three times slower than the
     * single-GPU one), no spare buffer to clear -- it re-zeroes each row it has read, and the next writes to
     * it come after the step's second barrier.  `zero` is then ignored.
     * Every buffer must then be n * 12 + 32 floats long (zero before the first step): row 0 belongs to rank 0
     * alone, so a rank's partial loss goes into word n * 12 + 16 * (step & 1) of its OWN buffer instead of row
     * 0's pad; the update kernels sum that word over `peers` and clear the other parity's for the next step. */
    int push_per;
} gs2d_grad_exchange;
/* Optional extras of a step: the learnable colour weight and the optimizer bookkeeping that the
 * reference's prune / add control needs (GaussianSplats_Represent.py:31-36 rgb_W, :62-67
 * colours = _features_dc * rgb_W, :98-172 removal_control / adaptive_control).
 *   rgb_w [n]            learnable per-Gaussian colour weight; NULL: weight 1, not a parameter
 *   adan_state_w [4][n]  its Adan state (exp_avg, exp_avg_sq, exp_avg_diff, previous gradient)
 *   rgb_w_grad_out [n]   parity tests: d loss / d rgb_w (may be NULL)
 *   adan_step            the optimizer's group['step'] used for the bias corrections
 *                        (optimizer.py:165-172); 0: use `step`.  It differs from the iteration
 *                        number once update_optimizer() rebuilt the optimizer (:93-97)
 *   adan_fresh           != 0: the optimizer state was (re)created since the last update, so the
 *                        previous gradient := this gradient (optimizer.py:183-185; the state
 *                        planes must have been zeroed by the caller).  Ignored when adan_step == 0
 *                        (then `step == 1` decides, as gs2d_fit_step does)
 *   skip_update          != 0: the UPDATE phase reports the loss and re-zeroes the gradient rows
 *                        but leaves parameters and optimizer state untouched and does not
 *                        pre-count the next step -- the reference's iterations in which the
 *                        control replaced the parameter tensors after backward(), so that
 *                        optimizer.step() saw no gradients (train_iter :199-204)
 *   slots                != 0: fixed-capacity tile slots instead of tile ranges -- tile t owns records
 *                        [t*256, t*256+256) of the packed region (256 = all the reference renders
 *                        of a tile, forward.cu:552), so nothing has to be counted or scanned before
 *                        the scatter and the UPDATE phase writes the next step's records itself: a
 *                        steady-state iteration is TWO kernels.  Needs cap >= tiles*256 and the same
 *                        value on every call of a workspace; `precounted` then means "the previous
 *                        UPDATE already scattered these parameters".  A tile that would take a 257th
 *                        record makes the step a no-op like a cap overflow, reported as a NEGATIVE
 *                        overflow_step: re-initialise the workspace and continue with slots = 0
 *                        (only ranges can tell WHICH 256 ids a fuller tile renders).  ids_sorted_out
 *                        is then indexed by slot (t*256 + rank).  Pairs whose pixel rectangle inside the
 *                        tile is empty (no pixel of the tile can pass the alpha test: they add nothing to
 *                        the image and get no gradient) are counted -- num_intersects, max_per_tile and the
 *                        257th-pair rule see every pair, like the reference -- but take no slot: the tile
 *                        kernel walks a quarter fewer records, same results.  Only a call that scatters
 *                        itself (precounted == 0) AND passes ids_sorted_out stores them all, so that the
 *                        list it returns is the reference's. */
typedef struct {
    float* rgb_w;
    float* adan_state_w;
    float* rgb_w_grad_out;
    int adan_step;
    int adan_fresh;
    int skip_update;
    int slots;
    int all_per_tile;    /* != 0: every Gaussian of a tile is rendered and gets its gradient, in batches of 256,
                          * where the reference stops after the first 256 (forward.cu:566-614, SURVEY Q1 /
                          * 8f.4).  Needs slots == 0 (tile ranges).  Off = parity with the reference. */
    int analytic_grad;   /* != 0: the projection backward uses the analytic gradient of the forward instead
                          * of the reference's formula (see
                          * gs2d_project_chol_bwd_analytic).  Off = parity with the reference. */
    float* params8;      /* != NULL: the three parameter tensors are the columns 0-1 / 2-4 / 5-7 of this ONE
                          * fp32 [n][8] buffer (16-byte aligned; rows = xyz 2 | chol 3 | features 3), and the
                          * xyz_raw / chol_raw / features arguments must be params8, params8 + 2, params8 + 5.
                          * Every per-Gaussian kernel then moves a Gaussian's parameters as two aligned 16-byte
                          * accesses where the reference's AoS tensors (cast to float2 / float3 at
                          * bindings.cu:456-465) take eight 4-byte ones.  With a sharded update
                          * (gs2d_grad_exchange) xyz_peers then lists the ranks' params8 buffers and
                          * chol_peers / feat_peers are unused. */
} gs2d_fit_options;
int gs2d_fit_step_band(int n, int cap, int img_h, int img_w, float* xyz_raw, float* chol_raw, float* features,
                       float* adan_state, const void* target, int target_format, void* work, size_t work_bytes,
                       int step, float lr, float beta1, float beta2, float beta3, float eps, float weight_decay,
                       int precounted, int tile_row_begin, int tile_row_end, int phases,
                       const gs2d_grad_exchange* exchange, const gs2d_fit_options* options, float* out_img,
                       float* param_grads_out, int32_t* ids_sorted_out, gs2d_fit_report* report_host,
                       gs2d_stream_t stream);
/* The Adan step of the operator loop as one kernel over all parameter tensors (replaces the ~20
 * torch._foreach_* launches of _multi_tensor_adan, optimizer.py:296-362; defaults of :65-74;
 * the reference anticipates an external fused kernel at :365-395).  Per element, g the gradient:
 *   diff = g - prev (0 when `first`);  m = b1 m + (1-b1) g;  d = b2 d + (1-b2) diff;
 *   u = g + b2 diff;  n = b3 n + (1-b3) u^2;  denom = sqrt(n)/sqrt(1-b3^step) + eps;
 *   p = (p - lr/(1-b1^step) m/denom - lr b2/(1-b2^step) d/denom) / (1 + lr wd);  prev = g
 * The six tables are HOST arrays of num_tensors DEVICE pointers (fp32, contiguous, counts[i]
 * elements each); `first` != 0 at the optimizer's first step (optimizer.py:183-185).
 * step_dev (DEVICE int32, may be NULL): when set, the step count is read from there at run time instead
 * of `step` -- a captured CUDA graph replays the same arguments every iteration, so the caller keeps the
 * counter on the device and bumps it before each step. */
int gs2d_adan_step(int num_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                   float* const* exp_avg_sq, float* const* exp_avg_diff, float* const* prev_grad, const int* counts,
                   float lr, float beta1, float beta2, float beta3, float eps, float weight_decay, int step,
                   const int* step_dev, int first, gs2d_stream_t stream);
/* The reference's prune (GaussianSplats_Represent.py:98-128, :147-171): drop the `remove_count`
 * Gaussians with the smallest |rgb_w| (torch.norm over dim 1 of an [N,1] tensor, then torch.sort,
 * :103-104 / :146-147; ties -- which torch.sort leaves unspecified -- go lowest index first) and
 * compact the four parameter tensors in index order (boolean-mask indexing, :111-114).  Built
 * from this library's own onesweep sort (32 key bits) and look-back scan plus two small kernels;
 * sync-free.  Outputs hold n - remove_count rows and must not alias the inputs.  keep_index_out
 * [n - remove_count] (may be NULL) receives the surviving rows' old indices.
 * scratch: gs2d_prune_scratch_bytes(n). */
size_t gs2d_prune_scratch_bytes(int n);
int gs2d_prune_lowest(int n, int remove_count, const float* rgb_w, const float* xyz_raw, const float* chol_raw,
                      const float* features, float* rgb_w_out, float* xyz_out, float* chol_out, float* features_out,
                      int32_t* keep_index_out, void* scratch, size_t scratch_bytes, gs2d_stream_t stream);
/* Decode side: the quantised wire format of the codec, rendered without ever
 * materialising fp32 Gaussians.  Same result as the reference's forward_quantize in eval mode
 * (GaussianSplats_Compress.py:71-84) fed with the same codes:
 *   xyz_half     uint16 [n][2]  IEEE fp16 bits of _xyz (FakeQuantizationHalf, quantize.py:15-24);
 *                               means2d = tanh(float(xyz_half))
 *   chol_code    uint8  [n][3]  6-bit codes 0..63 (UniformQuantizer.compress, quantize.py:82-84);
 *                               L = code * chol_scale[c] + chol_beta[c] (+ the Cholesky bound)
 *   color_index  uint8  [n][2]  residual-VQ indices of layer 0 and 1, 0..7 each;
 *                               colour = codebooks[0][i0] + codebooks[1][i1]   (quantize.py:146-150)
 *   codebooks    fp32 [2][8][3] device memory
 * 9 bytes per Gaussian instead of 32.  Opacity is 1 (the model's `_opacity` buffer).  Entropy
 * decoding of the code streams (constriction ANS, quantize.py:152-180) stays on the CPU. */
typedef struct {
    const uint16_t* xyz_half;
    const uint8_t* chol_code;
    const uint8_t* color_index;
    const float* codebooks;
    float chol_scale[3];
    float chol_beta[3];
} gs2d_quantized_frame;
/* gs2d_render_chol on the wire format (work / outputs as there). */
int gs2d_render_quantized(int n, int cap, int img_h, int img_w, const gs2d_quantized_frame* frame, void* work,
                          size_t work_bytes, float* out_img, int out_chw, int out_clamp, int32_t* final_idx,
                          int32_t* ids_sorted_out, gs2d_stream_t stream);
/* The fp32 parameter tensors the wire format stands for (_xyz [n][2] before tanh, _cholesky [n][3]
 * before the bound, colours [n][3]): what a P-frame fit starts from after decoding frame i-1. */
int gs2d_dequantize(int n, const gs2d_quantized_frame* frame, float* xyz_raw, float* chol_raw, float* features,
                    gs2d_stream_t stream);
/* Frame pipe: raw frames from PINNED host memory into a ring of `depth` device slots
 * (device_slots: depth * frame_bytes bytes, caller-owned) on the pipe's own copy stream, so the
 * copy of the next frame runs under the current frame's fit step.  No counterpart in the
 * reference, which decodes all frames on the CPU up front (utils.py:134-156) and moves each to
 * the GPU synchronously (train_video_Represent.py:43).  Per frame:
 *   push(host_frame) -> slot     copy queued (after the slot's previous consumer released it)
 *   acquire(slot, stream)        `stream` waits for that copy; returns the slot's device pointer
 *   ... enqueue the work that reads the frame on `stream` (gs2d_fit_step, gs2d_i420_to_rgb) ...
 *   release(slot, stream)        the slot may be overwritten once the work queued so far is done
 * Nothing here blocks the host.  The host frame must stay valid until its copy ran. */
#define GS2D_PIPE_MAX_DEPTH 8
typedef struct gs2d_frame_pipe gs2d_frame_pipe;
int gs2d_frame_pipe_create(void* device_slots, size_t frame_bytes, int depth, gs2d_frame_pipe** out);
int gs2d_frame_pipe_destroy(gs2d_frame_pipe* pipe);
/* Optional: every pushed frame (of `target_format`: GS2D_TARGET_F32_CHW / _U8_HWC / _I420, img_h x img_w) is
 * also converted on the copy stream, right behind its copy, into the tile kernel's own order
 * (gs2d_fit_target_prepare) -- prepared_slots: depth * gs2d_fit_target_bytes(H, W) bytes, caller-owned --
 * and acquire() then hands out that copy, to be passed with GS2D_TARGET_TILED.  NULL switches it off. */
int gs2d_frame_pipe_set_prepare(gs2d_frame_pipe* pipe, int img_h, int img_w, int target_format, float* prepared_slots);
int gs2d_frame_pipe_push(gs2d_frame_pipe* pipe, const void* host_frame, int* slot_out);
int gs2d_frame_pipe_acquire(gs2d_frame_pipe* pipe, int slot, gs2d_stream_t stream, void** device_frame);
int gs2d_frame_pipe_release(gs2d_frame_pipe* pipe, int slot, gs2d_stream_t stream);
/* The reference's frame loader arithmetic, process_yuv_video (utils.py:134-156):
 * cv2.cvtColor(yuv, cv2.COLOR_YUV2RGB_I420) of one planar I420 frame -> uint8 RGB [H][W][3].
 * OpenCV's BT.601 limited-range 20-bit fixed-point formula, chroma of the pixel's 2x2 block
 * (oracle/yuv.py states it; bit-exact against cv2 over all 2^24 triples).  H and W even. */
int gs2d_i420_to_rgb(int img_h, int img_w, const uint8_t* yuv, uint8_t* rgb, gs2d_stream_t stream);
/* Render-only (decode): project_gaussians_2d + rasterize_gaussians_sum forward of the reference
 * (GaussianSplats_Represent.py:83-90) as three kernels, sync-free.  activate is a bit set:
 * GS2D_RENDER_ACTIVATE (1; any odd value): inputs are the raw parameters (tanh / bound applied here),
 * else means2d / L_elements as the projection API takes them; GS2D_RENDER_ALL_PER_TILE (4): every
 * Gaussian of a tile is rendered, not only the first 256 (see gs2d_fit_options.all_per_tile).  opacities may be NULL (= 1).  out_img fp32: planar [3][H][W] when out_chw,
 * else [H][W][3]; clamped to [0,1] when out_clamp.  final_idx [H][W] and ids_sorted_out may be
 * NULL.  work as above (gs2d_fit_init first; the Adan pointer may be NULL there). */
#define GS2D_RENDER_ACTIVATE 1
#define GS2D_RENDER_ALL_PER_TILE 4
#define GS2D_RENDER_PACKED8 8 /* means_or_xyz_raw is a 16-byte aligned fp32 [n][8] buffer of packed rows
                               * (gs2d_fit_options.params8), chol = its column 2; colors = its column 5 or
                               * a contiguous [n][3] tensor of its own */
int gs2d_render_chol(int n, int cap, int img_h, int img_w, const float* means_or_xyz_raw, const float* chol,
                     const float* colors, const float* opacities, int activate, void* work, size_t work_bytes,
                     float* out_img, int out_chw, int out_clamp, int32_t* final_idx, int32_t* ids_sorted_out,
                     gs2d_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GS2D_H */
