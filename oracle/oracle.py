"""numpy front-end of the C oracle (oracle/gs2d_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(gsvc_b200/) never imports this module.

Function names follow the reference's Python API (gsplat/gsplat/*.py in the
reference tree); each wraps the C function that restates the cited kernel.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgs2d_oracle.so")
_lib = None

_f = np.float32
_i = np.int32


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (recipe: oracle/Makefile)."""
    src = os.path.join(_HERE, "gs2d_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libgs2d_oracle.so"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.gs2d_oracle_cumsum.restype = ctypes.c_int64
        _lib.gs2d_oracle_num_threads.restype = ctypes.c_int
    return _lib


def _p(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(np.asarray(a), dtype=dt)


def num_threads() -> int:
    return int(lib().gs2d_oracle_num_threads())


def set_num_threads(n: int) -> None:
    lib().gs2d_oracle_set_num_threads(int(n))


def tile_bounds_of(H: int, W: int):
    return ((W + 15) // 16, (H + 15) // 16, 1)


# --- helpers.cuh:45-68 -----------------------------------------------------------
def compute_cov2d_bounds(cov2d):
    cov2d = _c(cov2d, _f)
    n = cov2d.shape[0]
    conics = np.zeros((n, 3), _f)
    radii = np.zeros((n, 1), _f)
    lib().gs2d_oracle_cov2d_bounds(n, _p(cov2d), _p(conics), _p(radii))
    return conics, radii


# --- helpers.cuh:27-43 -----------------------------------------------------------
def get_tile_bbox(xys, radii, tile_bounds):
    xys = _c(xys, _f)
    radii_f = _c(radii, _f)
    n = xys.shape[0]
    mn = np.zeros((n, 2), _i)
    mx = np.zeros((n, 2), _i)
    lib().gs2d_oracle_tile_bbox(n, _p(xys), _p(radii_f), int(tile_bounds[0]), int(tile_bounds[1]), _p(mn), _p(mx))
    return mn, mx


# --- foward2d.cu:12-69 -----------------------------------------------------------
def project_gaussians_2d_forward(means2d, L_elements, H, W, tile_bounds):
    means2d = _c(means2d, _f)
    L = _c(L_elements, _f)
    n = means2d.shape[0]
    xys = np.zeros((n, 2), _f)
    depths = np.zeros((n,), _f)
    radii = np.zeros((n,), _i)
    conics = np.zeros((n, 3), _f)
    nth = np.zeros((n,), _i)
    lib().gs2d_oracle_project_chol_fwd(n, _p(means2d), _p(L), int(H), int(W), int(tile_bounds[0]),
                                       int(tile_bounds[1]), _p(xys), _p(depths), _p(radii), _p(conics), _p(nth))
    return xys, depths, radii, conics, nth


# --- foward2d.cu:71-124 ----------------------------------------------------------
def project_gaussians_2d_scale_rot_forward(means2d, scales2d, rotation, H, W, tile_bounds):
    means2d = _c(means2d, _f)
    scales2d = _c(scales2d, _f)
    rotation = _c(rotation, _f).reshape(-1)
    n = means2d.shape[0]
    xys = np.zeros((n, 2), _f)
    depths = np.zeros((n,), _f)
    radii = np.zeros((n,), _i)
    conics = np.zeros((n, 3), _f)
    nth = np.zeros((n,), _i)
    lib().gs2d_oracle_project_scalerot_fwd(n, _p(means2d), _p(scales2d), _p(rotation), int(H), int(W),
                                           int(tile_bounds[0]), int(tile_bounds[1]), _p(xys), _p(depths),
                                           _p(radii), _p(conics), _p(nth))
    return xys, depths, radii, conics, nth


# --- gsplat/gsplat/utils.py:99-118 -----------------------------------------------
def compute_cumulative_intersects(num_tiles_hit):
    nth = _c(num_tiles_hit, _i)
    cum = np.zeros_like(nth)
    total = lib().gs2d_oracle_cumsum(nth.shape[0], _p(nth), _p(cum))
    return int(total), cum


# --- forward.cu:100-136 ----------------------------------------------------------
def map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds):
    xys = _c(xys, _f)
    depths = _c(depths, _f)
    radii = _c(radii, _i)
    cum = _c(cum_tiles_hit, _i)
    isect = np.zeros((num_intersects,), np.int64)
    gids = np.zeros((num_intersects,), _i)
    lib().gs2d_oracle_map_intersects(int(num_points), _p(xys), _p(depths), _p(radii), _p(cum),
                                     int(tile_bounds[0]), int(tile_bounds[1]), _p(isect), _p(gids))
    return isect, gids


# --- gsplat/gsplat/utils.py:164-165 ----------------------------------------------
def sort_by_key(isect_ids, gaussian_ids):
    k = _c(isect_ids, np.int64)
    v = _c(gaussian_ids, _i)
    ko = np.zeros_like(k)
    vo = np.zeros_like(v)
    lib().gs2d_oracle_sort(k.shape[0], _p(k), _p(v), _p(ko), _p(vo))
    return ko, vo


# --- forward.cu:141-163 ----------------------------------------------------------
def get_tile_bin_edges(num_intersects, isect_ids_sorted, rows=None):
    k = _c(isect_ids_sorted, np.int64)
    rows = int(num_intersects if rows is None else rows)
    bins = np.zeros((rows, 2), _i)
    lib().gs2d_oracle_tile_bin_edges(int(num_intersects), _p(k), _p(bins))
    return bins


# --- gsplat/gsplat/utils.py:121-167 ----------------------------------------------
def bin_and_sort_gaussians(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds):
    isect, gids = map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds)
    isect_s, gids_s = sort_by_key(isect, gids)
    rows = max(int(num_intersects), int(tile_bounds[0]) * int(tile_bounds[1]))
    bins = get_tile_bin_edges(num_intersects, isect_s, rows=rows)
    return isect, gids, isect_s, gids_s, bins


# --- forward.cu:512-627 ----------------------------------------------------------
def rasterize_sum_forward(H, W, gaussian_ids_sorted, tile_bins, xys, conics, colors, opacity):
    tb = tile_bounds_of(H, W)
    ids = _c(gaussian_ids_sorted, _i)
    bins = _c(tile_bins, _i)
    assert bins.shape[0] >= tb[0] * tb[1], "tile_bins must cover every tile (SURVEY Q6)"
    xys = _c(xys, _f)
    conics = _c(conics, _f)
    colors = _c(colors, _f)
    opacity = _c(opacity, _f).reshape(-1)
    out = np.zeros((H, W, 3), _f)
    Ts = np.zeros((H, W), _f)
    fidx = np.zeros((H, W), _i)
    lib().gs2d_oracle_rasterize_sum_fwd(int(H), int(W), tb[0], tb[1], _p(ids), _p(bins), _p(xys), _p(conics),
                                        _p(colors), _p(opacity), _p(out), _p(Ts), _p(fidx))
    return out, Ts, fidx


# --- backward.cu:696-862 ---------------------------------------------------------
def rasterize_sum_backward(H, W, gaussian_ids_sorted, tile_bins, xys, conics, colors, opacity, final_idx, v_out):
    tb = tile_bounds_of(H, W)
    ids = _c(gaussian_ids_sorted, _i)
    bins = _c(tile_bins, _i)
    xys = _c(xys, _f)
    conics = _c(conics, _f)
    colors = _c(colors, _f)
    opacity = _c(opacity, _f).reshape(-1)
    fidx = _c(final_idx, _i)
    v_out = _c(v_out, _f)
    n = xys.shape[0]
    v_xy = np.zeros((n, 2), _f)
    v_conic = np.zeros((n, 3), _f)
    v_rgb = np.zeros((n, 3), _f)
    v_op = np.zeros((n, 1), _f)
    lib().gs2d_oracle_rasterize_sum_bwd(n, int(H), int(W), tb[0], tb[1], _p(ids), _p(bins), _p(xys), _p(conics),
                                        _p(colors), _p(opacity), _p(fidx), _p(v_out), _p(v_xy), _p(v_conic),
                                        _p(v_rgb), _p(v_op))
    return v_xy, v_conic, v_rgb, v_op


# --- backward2d.cu:8-51 ----------------------------------------------------------
def project_gaussians_2d_backward(L_elements, H, W, radii, conics, v_xy, v_conic):
    L = _c(L_elements, _f)
    radii = _c(radii, _i)
    conics = _c(conics, _f)
    v_xy = _c(v_xy, _f)
    v_conic = _c(v_conic, _f)
    n = L.shape[0]
    v_cov2d = np.zeros((n, 3), _f)
    v_mean = np.zeros((n, 2), _f)
    v_L = np.zeros((n, 3), _f)
    lib().gs2d_oracle_project_chol_bwd(n, _p(L), int(H), int(W), _p(radii), _p(conics), _p(v_xy), _p(v_conic),
                                       _p(v_cov2d), _p(v_mean), _p(v_L))
    return v_cov2d, v_mean, v_L


# --- backward2d.cu:53-110 --------------------------------------------------------
def project_gaussians_2d_scale_rot_backward(scales2d, rotation, H, W, radii, conics, v_xy, v_conic):
    scales2d = _c(scales2d, _f)
    rotation = _c(rotation, _f).reshape(-1)
    radii = _c(radii, _i)
    conics = _c(conics, _f)
    v_xy = _c(v_xy, _f)
    v_conic = _c(v_conic, _f)
    n = scales2d.shape[0]
    v_cov2d = np.zeros((n, 3), _f)
    v_mean = np.zeros((n, 2), _f)
    v_scale = np.zeros((n, 2), _f)
    v_rot = np.zeros((n, 1), _f)
    lib().gs2d_oracle_project_scalerot_bwd(n, _p(scales2d), _p(rotation), int(H), int(W), _p(radii), _p(conics),
                                           _p(v_xy), _p(v_conic), _p(v_cov2d), _p(v_mean), _p(v_scale), _p(v_rot))
    return v_cov2d, v_mean, v_scale, v_rot


# --- the whole path, as GaussianSplats_Represent.py:83-90 drives it ---------------
def render(means2d, L_elements, colors, opacity, H, W):
    """project (Cholesky) -> bin -> sum-rasterise forward.  Returns a dict of every
    intermediate so tests can compare stage by stage."""
    tb = tile_bounds_of(H, W)
    xys, depths, radii, conics, nth = project_gaussians_2d_forward(means2d, L_elements, H, W, tb)
    I, cum = compute_cumulative_intersects(nth)
    n = xys.shape[0]
    if I < 1:
        return dict(xys=xys, depths=depths, radii=radii, conics=conics, num_tiles_hit=nth, num_intersects=0,
                    out_img=np.ones((H, W, 3), _f), final_idx=np.zeros((H, W), _i))
    isect, gids, isect_s, gids_s, bins = bin_and_sort_gaussians(n, I, xys, depths, radii, cum, tb)
    out, Ts, fidx = rasterize_sum_forward(H, W, gids_s, bins, xys, conics, colors, opacity)
    return dict(xys=xys, depths=depths, radii=radii, conics=conics, num_tiles_hit=nth, cum_tiles_hit=cum,
                num_intersects=I, isect_ids=isect, gaussian_ids=gids, isect_ids_sorted=isect_s,
                gaussian_ids_sorted=gids_s, tile_bins=bins, out_img=out, final_Ts=Ts, final_idx=fidx)


def fit_grads(means2d, L_elements, colors, opacity, H, W, v_out_fn):
    """forward + backward of the whole path; `v_out_fn(out_img)` returns dLoss/d out_img."""
    r = render(means2d, L_elements, colors, opacity, H, W)
    v_out = _c(v_out_fn(r["out_img"]), _f)
    if r["num_intersects"] < 1:
        n = r["xys"].shape[0]
        r.update(v_xy=np.zeros((n, 2), _f), v_conic=np.zeros((n, 3), _f), v_colors=np.zeros((n, 3), _f),
                 v_opacity=np.zeros((n, 1), _f), v_mean2d=np.zeros((n, 2), _f), v_L=np.zeros((n, 3), _f))
        return r
    v_xy, v_conic, v_rgb, v_op = rasterize_sum_backward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], r["xys"],
                                                        r["conics"], colors, opacity, r["final_idx"], v_out)
    v_cov2d, v_mean, v_L = project_gaussians_2d_backward(L_elements, H, W, r["radii"], r["conics"], v_xy, v_conic)
    r.update(v_out=v_out, v_xy=v_xy, v_conic=v_conic, v_colors=v_rgb, v_opacity=v_op, v_cov2d=v_cov2d,
             v_mean2d=v_mean, v_L=v_L)
    return r
