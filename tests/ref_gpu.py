"""torch front-end of oracle/_ref/libgsplat_ref.so -- the REFERENCE's own CUDA kernels, compiled
from /root/reference by oracle/build_ref.sh (test infrastructure; GPU box only).

`pipeline()` strings them together exactly as the reference's Python does
(gsplat/gsplat/rasterize_sum.py:93-186 + utils.py:99-167): torch.cumsum, .item(), torch.sort,
torch.gather in between the kernels."""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(os.path.dirname(_HERE), "oracle", "_ref", "libgsplat_ref.so")
_lib = None


def available() -> bool:
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(REF_LIB)
    return _lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _s():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ck(rc, what):
    if rc != 0:
        raise RuntimeError(f"reference launch {what} failed: cudaError {rc}")


f32, i32, i64 = torch.float32, torch.int32, torch.int64


def compute_cov2d_bounds(cov2d):
    n = cov2d.shape[0]
    conics = torch.empty((n, 3), dtype=f32, device=cov2d.device)
    radii = torch.empty((n, 1), dtype=f32, device=cov2d.device)
    _ck(lib().ref_compute_cov2d_bounds(n, _p(cov2d.contiguous()), _p(conics), _p(radii), _s()), "cov2d_bounds")
    return conics, radii


def _proj_out(n, dev):
    return (torch.empty((n, 2), dtype=f32, device=dev), torch.empty((n,), dtype=f32, device=dev),
            torch.empty((n,), dtype=i32, device=dev), torch.empty((n, 3), dtype=f32, device=dev),
            torch.empty((n,), dtype=i32, device=dev))


def project_gaussians_2d_forward(means2d, L, H, W, tile_bounds, clip=0.01):
    n = means2d.shape[0]
    xys, depths, radii, conics, nth = _proj_out(n, means2d.device)
    _ck(lib().ref_project_gaussians_2d_forward(
        n, _p(means2d.contiguous()), _p(L.contiguous()), H, W, tile_bounds[0], tile_bounds[1], ctypes.c_float(clip),
        _p(xys), _p(depths), _p(radii), _p(conics), _p(nth), _s()), "project_fwd")
    return xys, depths, radii, conics, nth


def project_gaussians_2d_scale_rot_forward(means2d, scales, rot, H, W, tile_bounds, clip=0.01):
    n = means2d.shape[0]
    xys, depths, radii, conics, nth = _proj_out(n, means2d.device)
    _ck(lib().ref_project_gaussians_2d_scale_rot_forward(
        n, _p(means2d.contiguous()), _p(scales.contiguous()), _p(rot.contiguous()), H, W, tile_bounds[0],
        tile_bounds[1], ctypes.c_float(clip), _p(xys), _p(depths), _p(radii), _p(conics), _p(nth), _s()),
        "project_sr_fwd")
    return xys, depths, radii, conics, nth


def project_gaussians_2d_backward(means2d, L, H, W, radii, conics, v_xy, v_conic):
    n = means2d.shape[0]
    dev = means2d.device
    v_cov2d = torch.empty((n, 3), dtype=f32, device=dev)
    v_mean = torch.empty((n, 2), dtype=f32, device=dev)
    v_L = torch.empty((n, 3), dtype=f32, device=dev)
    v_depth = torch.zeros((n,), dtype=f32, device=dev)
    _ck(lib().ref_project_gaussians_2d_backward(
        n, _p(means2d.contiguous()), _p(L.contiguous()), H, W, _p(radii), _p(conics), _p(v_xy.contiguous()),
        _p(v_depth), _p(v_conic.contiguous()), _p(v_cov2d), _p(v_mean), _p(v_L), _s()), "project_bwd")
    return v_cov2d, v_mean, v_L


def project_gaussians_2d_scale_rot_backward(means2d, scales, rot, H, W, radii, conics, v_xy, v_conic):
    n = means2d.shape[0]
    dev = means2d.device
    v_cov2d = torch.empty((n, 3), dtype=f32, device=dev)
    v_mean = torch.empty((n, 2), dtype=f32, device=dev)
    v_scale = torch.empty((n, 2), dtype=f32, device=dev)
    v_rot = torch.empty((n, 1), dtype=f32, device=dev)
    v_depth = torch.zeros((n,), dtype=f32, device=dev)
    _ck(lib().ref_project_gaussians_2d_scale_rot_backward(
        n, _p(means2d.contiguous()), _p(scales.contiguous()), _p(rot.contiguous()), H, W, _p(radii), _p(conics),
        _p(v_xy.contiguous()), _p(v_depth), _p(v_conic.contiguous()), _p(v_cov2d), _p(v_mean), _p(v_scale), _p(v_rot),
        _s()), "project_sr_bwd")
    return v_cov2d, v_mean, v_scale, v_rot


def map_gaussian_to_intersects(n, I, xys, depths, radii, cum, tile_bounds):
    isect = torch.empty((I,), dtype=i64, device=xys.device)
    gids = torch.empty((I,), dtype=i32, device=xys.device)
    _ck(lib().ref_map_gaussian_to_intersects(n, I, _p(xys), _p(depths), _p(radii), _p(cum), tile_bounds[0],
                                             tile_bounds[1], _p(isect), _p(gids), _s()), "map")
    return isect, gids


def get_tile_bin_edges(I, isect_sorted, rows):
    bins = torch.empty((rows, 2), dtype=i32, device=isect_sorted.device)
    _ck(lib().ref_get_tile_bin_edges(I, _p(isect_sorted), _p(bins), rows, _s()), "edges")
    return bins


def rasterize_sum_forward(H, W, ids_sorted, tile_bins, xys, conics, colors, opac, background=None):
    dev = xys.device
    if background is None:
        background = torch.ones(3, dtype=f32, device=dev)
    out = torch.empty((H, W, 3), dtype=f32, device=dev)
    Ts = torch.empty((H, W), dtype=f32, device=dev)
    fidx = torch.empty((H, W), dtype=i32, device=dev)
    _ck(lib().ref_rasterize_sum_forward(H, W, _p(ids_sorted), _p(tile_bins), _p(xys), _p(conics), _p(colors),
                                        _p(opac), _p(background), _p(out), _p(Ts), _p(fidx), _s()), "raster_fwd")
    return out, Ts, fidx


def rasterize_sum_backward(H, W, ids_sorted, tile_bins, xys, conics, colors, opac, final_Ts, final_idx, v_out,
                           background=None):
    dev = xys.device
    n = xys.shape[0]
    if background is None:
        background = torch.ones(3, dtype=f32, device=dev)
    v_alpha = torch.zeros((H, W), dtype=f32, device=dev)
    v_xy = torch.empty((n, 2), dtype=f32, device=dev)
    v_conic = torch.empty((n, 3), dtype=f32, device=dev)
    v_col = torch.empty((n, 3), dtype=f32, device=dev)
    v_op = torch.empty((n, 1), dtype=f32, device=dev)
    _ck(lib().ref_rasterize_sum_backward(n, H, W, _p(ids_sorted), _p(tile_bins), _p(xys), _p(conics), _p(colors),
                                         _p(opac), _p(background), _p(final_Ts), _p(final_idx), _p(v_out.contiguous()),
                                         _p(v_alpha), _p(v_xy), _p(v_conic), _p(v_col), _p(v_op), _s()), "raster_bwd")
    return v_xy, v_conic, v_col, v_op


def nd_rasterize_sum_forward(H, W, ids_sorted, tile_bins, xys, conics, colors, opac):
    """the reference's unbound N-channel forward (forward.cu:630-710); colors [n, C]"""
    dev = xys.device
    C = colors.shape[1]
    background = torch.ones(C, dtype=f32, device=dev)
    out = torch.empty((H, W, C), dtype=f32, device=dev)
    Ts = torch.empty((H, W), dtype=f32, device=dev)
    fidx = torch.empty((H, W), dtype=i32, device=dev)
    _ck(lib().ref_nd_rasterize_sum_forward(H, W, C, _p(ids_sorted), _p(tile_bins), _p(xys), _p(conics),
                                           _p(colors.contiguous()), _p(opac), _p(background), _p(out), _p(Ts),
                                           _p(fidx), _s()), "nd_raster_fwd")
    return out, Ts, fidx


def nd_rasterize_sum_backward(H, W, ids_sorted, tile_bins, xys, conics, colors, opac, final_Ts, final_idx, v_out):
    """the reference's unbound N-channel backward (backward.cu:864-978).  It walks sorted indices
    final_idx - 1 down to the tile's first, so pass final_idx = (one past the last index that is to
    receive a gradient)."""
    dev = xys.device
    n, C = xys.shape[0], colors.shape[1]
    background = torch.ones(C, dtype=f32, device=dev)
    v_alpha = torch.zeros((H, W), dtype=f32, device=dev)
    v_xy = torch.empty((n, 2), dtype=f32, device=dev)
    v_conic = torch.empty((n, 3), dtype=f32, device=dev)
    v_col = torch.empty((n, C), dtype=f32, device=dev)
    v_op = torch.empty((n, 1), dtype=f32, device=dev)
    ws = torch.empty((H, W, C), dtype=f32, device=dev)
    _ck(lib().ref_nd_rasterize_sum_backward(n, H, W, C, _p(ids_sorted), _p(tile_bins), _p(xys), _p(conics),
                                            _p(colors.contiguous()), _p(opac), _p(background), _p(final_Ts),
                                            _p(final_idx.contiguous()), _p(v_out.contiguous()), _p(v_alpha), _p(v_xy),
                                            _p(v_conic), _p(v_col), _p(v_op), _p(ws), _s()), "nd_raster_bwd")
    return v_xy, v_conic, v_col, v_op


def bin_and_sort(n, xys, depths, radii, nth, tile_bounds, stable=True):
    """utils.py:99-167 of the reference around its kernels.  tile_bins gets max(I, tiles) rows so
    the reference kernel stays in bounds."""
    cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
    I = int(cum[-1].item())
    if I < 1:
        return dict(num_intersects=0, cum=cum)
    isect, gids = map_gaussian_to_intersects(n, I, xys, depths, radii, cum, tile_bounds)
    isect_sorted, order = torch.sort(isect, stable=stable)
    gids_sorted = torch.gather(gids, 0, order)
    rows = max(I, tile_bounds[0] * tile_bounds[1])
    bins = get_tile_bin_edges(I, isect_sorted, rows)
    return dict(num_intersects=I, cum=cum, isect_ids=isect, gaussian_ids=gids, isect_ids_sorted=isect_sorted,
                gaussian_ids_sorted=gids_sorted, tile_bins=bins)


def pipeline(means2d, L, colors, opac, H, W, v_out_fn=None, stable=True):
    """project (Cholesky) -> bin -> raster forward [-> raster backward -> project backward]."""
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    n = means2d.shape[0]
    xys, depths, radii, conics, nth = project_gaussians_2d_forward(means2d, L, H, W, tb)
    r = dict(xys=xys, depths=depths, radii=radii, conics=conics, num_tiles_hit=nth)
    r.update(bin_and_sort(n, xys, depths, radii, nth, tb, stable=stable))
    if r["num_intersects"] < 1:
        return r
    out, Ts, fidx = rasterize_sum_forward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], xys, conics, colors, opac)
    r.update(out_img=out, final_Ts=Ts, final_idx=fidx)
    if v_out_fn is not None:
        v_out = v_out_fn(out).contiguous()
        v_xy, v_conic, v_col, v_op = rasterize_sum_backward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], xys,
                                                            conics, colors, opac, Ts, fidx, v_out)
        v_cov2d, v_mean, v_L = project_gaussians_2d_backward(means2d, L, H, W, radii, conics, v_xy, v_conic)
        r.update(v_out=v_out, v_xy=v_xy, v_conic=v_conic, v_colors=v_col, v_opacity=v_op, v_mean2d=v_mean, v_L=v_L)
    return r
