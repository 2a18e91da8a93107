"""`gsplat.cuda` mirror: the reference's pybind functions on the hot path (ext.cpp:10-23 of
the reference), same names, same argument order and return tuples, implemented as thin torch
shims over the C ABI of libgs2d.so.  Tensors must be CUDA / contiguous like the reference's
CHECK_INPUT (bindings.h:9-14); outputs are fresh tensors on the input's device; launches go to
torch's current stream under a device guard (the reference used the legacy stream).
"""
from __future__ import annotations

from typing import Tuple

import torch

from .. import _lib

_f32, _i32, _i64 = torch.float32, torch.int32, torch.int64


def _chk(t: torch.Tensor, name: str, dtype=None) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


class _guard:
    """device guard that costs nothing when the tensor already lives on the current device"""

    __slots__ = ("idx", "prev")

    def __init__(self, dev: torch.device):
        self.idx = dev.index if dev.index is not None else torch.cuda.current_device()

    def __enter__(self):
        self.prev = torch.cuda.current_device()
        _lib.require_device(self.idx)
        if self.prev != self.idx:
            torch.cuda.set_device(self.idx)
        return self

    def __exit__(self, *exc):
        if self.prev != self.idx:
            torch.cuda.set_device(self.prev)
        return False


def _ptr(t):
    return None if t is None else t.data_ptr()


# --- bindings.cu:41-61 ---------------------------------------------------------------------
def compute_cov2d_bounds(num_pts: int, covs2d: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    _chk(covs2d, "covs2d", _f32)
    conics = torch.empty((num_pts, covs2d.size(1)), dtype=_f32, device=covs2d.device)
    radii = torch.empty((num_pts, 1), dtype=_f32, device=covs2d.device)
    with _guard(covs2d.device):
        _lib.check(_lib.load().gs2d_cov2d_bounds(num_pts, _ptr(covs2d), _ptr(conics), _ptr(radii),
                                                 _stream(covs2d.device)), "cov2d_bounds")
    return conics, radii


# --- bindings.cu:787-839 -------------------------------------------------------------------
def project_gaussians_2d_forward(num_points, means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh):
    _chk(means2d, "means2d", _f32)
    _chk(L_elements, "L_elements", _f32)
    dev = means2d.device
    xys = torch.empty((num_points, 2), dtype=_f32, device=dev)
    depths = torch.empty((num_points,), dtype=_f32, device=dev)
    radii = torch.empty((num_points,), dtype=_i32, device=dev)
    conics = torch.empty((num_points, 3), dtype=_f32, device=dev)
    nth = torch.empty((num_points,), dtype=_i32, device=dev)
    with _guard(dev):
        _lib.check(_lib.load().gs2d_project_chol_fwd(
            num_points, _ptr(means2d), _ptr(L_elements), int(img_height), int(img_width), int(tile_bounds[0]),
            int(tile_bounds[1]), float(clip_thresh), _ptr(xys), _ptr(depths), _ptr(radii), _ptr(conics), _ptr(nth),
            _stream(dev)), "project_chol_fwd")
    return xys, depths, radii, conics, nth


# --- bindings.cu:841-899 -------------------------------------------------------------------
def project_gaussians_2d_scale_rot_forward(num_points, means2d, scales2d, rotation, img_height, img_width,
                                           tile_bounds, clip_thresh):
    _chk(means2d, "means2d", _f32)
    _chk(scales2d, "scales2d", _f32)
    _chk(rotation, "rotation", _f32)
    dev = means2d.device
    xys = torch.empty((num_points, 2), dtype=_f32, device=dev)
    depths = torch.empty((num_points,), dtype=_f32, device=dev)
    radii = torch.empty((num_points,), dtype=_i32, device=dev)
    conics = torch.empty((num_points, 3), dtype=_f32, device=dev)
    nth = torch.empty((num_points,), dtype=_i32, device=dev)
    with _guard(dev):
        _lib.check(_lib.load().gs2d_project_scalerot_fwd(
            num_points, _ptr(means2d), _ptr(scales2d), _ptr(rotation), int(img_height), int(img_width),
            int(tile_bounds[0]), int(tile_bounds[1]), float(clip_thresh), _ptr(xys), _ptr(depths), _ptr(radii),
            _ptr(conics), _ptr(nth), _stream(dev)), "project_scalerot_fwd")
    return xys, depths, radii, conics, nth


def _rows(t: torch.Tensor, width: int, name: str):
    """(tensor, leading dimension) of a [N,width] fp32 gradient that may be a strided row view"""
    if t.dtype != _f32 or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA float32 tensor")
    if t.dim() != 2 or t.size(1) != width or t.stride(1) != 1 or t.stride(0) < width:
        t = t.contiguous().view(-1, width)
    return t, t.stride(0)


# --- bindings.cu:902-949 -------------------------------------------------------------------
def project_gaussians_2d_backward(num_points, means2d, L_elements, img_height, img_width, radii, conics, v_xy,
                                  v_depth, v_conic, analytic: bool = False):
    _chk(L_elements, "L_elements", _f32)
    _chk(radii, "radii", _i32)
    _chk(conics, "conics", _f32)
    dev = L_elements.device
    v_xy, ld_xy = _rows(v_xy, 2, "v_xy")
    v_conic, ld_conic = _rows(v_conic, 3, "v_conic")
    v_cov2d = torch.empty((num_points, 3), dtype=_f32, device=dev)
    v_mean2d = torch.empty((num_points, 2), dtype=_f32, device=dev)
    v_L = torch.empty((num_points, 3), dtype=_f32, device=dev)
    with _guard(dev):
        fn = _lib.load().gs2d_project_chol_bwd_analytic if analytic else _lib.load().gs2d_project_chol_bwd
        _lib.check(fn(
            num_points, _ptr(L_elements), int(img_height), int(img_width), _ptr(radii), _ptr(conics), _ptr(v_xy),
            ld_xy, _ptr(v_conic), ld_conic, _ptr(v_cov2d), _ptr(v_mean2d), _ptr(v_L), _stream(dev)),
            "project_chol_bwd")
    return v_cov2d, v_mean2d, v_L


# --- bindings.cu:951-1004 ------------------------------------------------------------------
def project_gaussians_2d_scale_rot_backward(num_points, means2d, scales2d, rotation, img_height, img_width, radii,
                                            conics, v_xy, v_depth, v_conic, analytic: bool = False):
    _chk(scales2d, "scales2d", _f32)
    _chk(rotation, "rotation", _f32)
    _chk(radii, "radii", _i32)
    _chk(conics, "conics", _f32)
    dev = scales2d.device
    v_xy, ld_xy = _rows(v_xy, 2, "v_xy")
    v_conic, ld_conic = _rows(v_conic, 3, "v_conic")
    v_cov2d = torch.empty((num_points, 3), dtype=_f32, device=dev)
    v_mean2d = torch.empty((num_points, 2), dtype=_f32, device=dev)
    v_scale = torch.empty((num_points, 2), dtype=_f32, device=dev)
    v_rot = torch.empty((num_points, 1), dtype=_f32, device=dev)
    with _guard(dev):
        fn = (_lib.load().gs2d_project_scalerot_bwd_analytic if analytic
              else _lib.load().gs2d_project_scalerot_bwd)
        _lib.check(fn(
            num_points, _ptr(scales2d), _ptr(rotation), int(img_height), int(img_width), _ptr(radii), _ptr(conics),
            _ptr(v_xy), ld_xy, _ptr(v_conic), ld_conic, _ptr(v_cov2d), _ptr(v_mean2d), _ptr(v_scale), _ptr(v_rot),
            _stream(dev)), "project_scalerot_bwd")
    return v_cov2d, v_mean2d, v_scale, v_rot


# --- torch.cumsum replacement (gsplat/gsplat/utils.py:116) -----------------------------------
def cumsum_tiles_hit(num_tiles_hit: torch.Tensor) -> torch.Tensor:
    _chk(num_tiles_hit, "num_tiles_hit", _i32)
    dev = num_tiles_hit.device
    n = num_tiles_hit.numel()
    cum = torch.empty((n,), dtype=_i32, device=dev)
    lib = _lib.load()
    nbytes = lib.gs2d_scan_scratch_bytes(n)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    with _guard(dev):
        _lib.check(lib.gs2d_cumsum_tiles_hit(n, _ptr(num_tiles_hit), _ptr(cum), None, None, _ptr(scratch), nbytes,
                                             _stream(dev)), "cumsum_tiles_hit")
    return cum


# --- bindings.cu:274-313 -------------------------------------------------------------------
def map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds):
    _chk(xys, "xys", _f32)
    _chk(depths, "depths", _f32)
    _chk(radii, "radii", _i32)
    _chk(cum_tiles_hit, "cum_tiles_hit", _i32)
    dev = xys.device
    isect_ids = torch.empty((num_intersects,), dtype=_i64, device=dev)
    gaussian_ids = torch.empty((num_intersects,), dtype=_i32, device=dev)
    with _guard(dev):
        _lib.check(_lib.load().gs2d_map_intersects(
            num_points, num_intersects, _ptr(xys), _ptr(depths), _ptr(radii), _ptr(cum_tiles_hit),
            int(tile_bounds[0]), int(tile_bounds[1]), _ptr(isect_ids), _ptr(gaussian_ids), _stream(dev)),
            "map_intersects")
    return isect_ids, gaussian_ids


# --- torch.sort + torch.gather replacement (gsplat/gsplat/utils.py:164-165) ------------------
def sort_pairs(isect_ids: torch.Tensor, gaussian_ids: torch.Tensor, begin_bit: int = 0, end_bit: int = 64):
    _chk(isect_ids, "isect_ids", _i64)
    _chk(gaussian_ids, "gaussian_ids", _i32)
    dev = isect_ids.device
    n = isect_ids.numel()
    keys_out = torch.empty_like(isect_ids)
    vals_out = torch.empty_like(gaussian_ids)
    lib = _lib.load()
    nbytes = lib.gs2d_sort_scratch_bytes(n)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    with _guard(dev):
        _lib.check(lib.gs2d_sort_pairs(n, _ptr(isect_ids), _ptr(gaussian_ids), _ptr(keys_out), _ptr(vals_out),
                                       int(begin_bit), int(end_bit), _ptr(scratch), nbytes, _stream(dev)),
                   "sort_pairs")
    return keys_out, vals_out


# --- bindings.cu:315-330 -------------------------------------------------------------------
def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: torch.Tensor, rows: int = None) -> torch.Tensor:
    """tile_bins has num_intersects rows like the reference unless `rows` is given;
    tile ids >= rows would be out of bounds in the reference and are a caller error here too."""
    _chk(isect_ids_sorted, "isect_ids_sorted", _i64)
    dev = isect_ids_sorted.device
    rows = int(num_intersects if rows is None else rows)
    tile_bins = torch.empty((rows, 2), dtype=_i32, device=dev)
    with _guard(dev):
        _lib.check(_lib.load().gs2d_tile_bin_edges(int(num_intersects), _ptr(isect_ids_sorted), _ptr(tile_bins), rows,
                                                   _stream(dev)), "tile_bin_edges")
    return tile_bins


# --- bindings.cu:400-469 -------------------------------------------------------------------
def rasterize_sum_forward(tile_bounds, block, img_size, gaussian_ids_sorted, tile_bins, xys, conics, colors,
                          opacities, background):
    for t, nm in ((gaussian_ids_sorted, "gaussian_ids_sorted"), (tile_bins, "tile_bins"), (xys, "xys"),
                  (conics, "conics"), (colors, "colors"), (opacities, "opacities"), (background, "background")):
        _chk(t, nm)
    if tuple(block)[:2] != (16, 16):
        raise RuntimeError("rasterize_sum_forward: only 16x16 blocks are supported (reference config.h:1-4)")
    if colors.size(1) != 3:
        raise RuntimeError("rasterize_sum_forward: only 3 channels are supported")
    W, H = int(img_size[0]), int(img_size[1])
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    if tile_bins.size(0) < tiles:
        raise RuntimeError(
            f"tile_bins has {tile_bins.size(0)} rows but the image has {tiles} tiles "
            "(the reference reads out of bounds here, SURVEY Q6)")
    dev = xys.device
    n, I = xys.size(0), gaussian_ids_sorted.numel()
    out_img = torch.empty((H, W, 3), dtype=_f32, device=dev)
    final_Ts = torch.empty((H, W), dtype=_f32, device=dev)
    final_idx = torch.empty((H, W), dtype=_i32, device=dev)
    lib = _lib.load()
    nbytes = lib.gs2d_raster_binned_scratch_bytes(I)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    with _guard(dev):
        _lib.check(lib.gs2d_rasterize_sum_fwd_binned(
            n, I, _ptr(gaussian_ids_sorted), _ptr(tile_bins), _ptr(xys), _ptr(conics), _ptr(colors), _ptr(opacities),
            H, W, _ptr(out_img), _ptr(final_Ts), _ptr(final_idx), _ptr(scratch), nbytes, _stream(dev)),
            "rasterize_sum_fwd_binned")
    return out_img, final_Ts, final_idx


# --- bindings.cu:706-779 -------------------------------------------------------------------
def rasterize_sum_backward(img_height, img_width, BLOCK_H, BLOCK_W, gaussian_ids_sorted, tile_bins, xys, conics,
                           colors, opacities, background, final_Ts, final_idx, v_output, v_output_alpha):
    _chk(xys, "xys", _f32)
    _chk(colors, "colors", _f32)
    if xys.dim() != 2 or xys.size(1) != 2:
        raise RuntimeError("xys must have dimensions (num_points, 2)")
    if colors.dim() != 2 or colors.size(1) != 3:
        raise RuntimeError("colors must have 2 dimensions")
    if (BLOCK_H, BLOCK_W) != (16, 16):
        raise RuntimeError("rasterize_sum_backward: only 16x16 blocks are supported")
    dev = xys.device
    H, W = int(img_height), int(img_width)
    n, I = xys.size(0), gaussian_ids_sorted.numel()
    v_output = v_output.contiguous()
    grads = torch.empty((n, 12), dtype=_f32, device=dev)
    lib = _lib.load()
    nbytes = lib.gs2d_raster_binned_scratch_bytes(I)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    with _guard(dev):
        _lib.check(lib.gs2d_rasterize_sum_bwd_binned(
            n, I, _ptr(gaussian_ids_sorted.contiguous()), _ptr(tile_bins.contiguous()), _ptr(xys),
            _ptr(conics.contiguous()), _ptr(colors), _ptr(opacities.contiguous()), H, W,
            _ptr(final_idx.contiguous()), _ptr(v_output), _ptr(grads), _ptr(scratch), nbytes, _stream(dev)),
            "rasterize_sum_bwd_binned")
    return grads[:, 0:2], grads[:, 2:5], grads[:, 5:8], grads[:, 8:9]
