"""Binning / sorting utilities with the reference's names, argument order and return tuples
(gsplat/gsplat/utils.py:12-167 of the reference), backed by libgs2d.so.  torch.cumsum /
torch.sort / torch.gather of the reference are replaced by the library's scan and onesweep
radix sort (same results: inclusive int32 prefix; stable ascending sort of the int64 keys)."""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import cuda as _C


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                               cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int]) -> Tuple[Tensor, Tensor]:
    """utils.py:12-50 -> (isect_ids int64 [I], gaussian_ids int32 [I])"""
    return _C.map_gaussian_to_intersects(num_points, num_intersects, xys.contiguous(), depths.contiguous(),
                                         radii.contiguous(), cum_tiles_hit.contiguous(), tile_bounds)


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor) -> Tensor:
    """utils.py:53-73 -> tile_bins int32 [num_intersects, 2] (reference shape)"""
    return _C.get_tile_bin_edges(num_intersects, isect_ids_sorted.contiguous())


def compute_cov2d_bounds(cov2d: Tensor) -> Tuple[Tensor, Tensor]:
    """utils.py:76-96 -> (conic [N,3], radius [N,1])"""
    assert cov2d.shape[-1] == 3, (
        f"Expected input cov2d to be of shape (*batch, 3) (upper triangular values), but got {tuple(cov2d.shape)}")
    num_pts = cov2d.shape[0]
    assert num_pts > 0
    return _C.compute_cov2d_bounds(num_pts, cov2d.contiguous())


def compute_cumulative_intersects(num_tiles_hit: Tensor) -> Tuple[int, Tensor]:
    """utils.py:99-118 -> (num_intersects: int, cum_tiles_hit int32 [N]).  Like the reference this
    returns a Python int and therefore waits for the device; the rasteriser's own
    path does not go through here."""
    cum_tiles_hit = _C.cumsum_tiles_hit(num_tiles_hit.contiguous().to(torch.int32))
    num_intersects = cum_tiles_hit[-1].item()
    return num_intersects, cum_tiles_hit


def bin_and_sort_gaussians(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                           cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int],
                           ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """utils.py:121-167 -> (isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins)"""
    isect_ids, gaussian_ids = map_gaussian_to_intersects(
        num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds)
    isect_ids_sorted, gaussian_ids_sorted = _C.sort_pairs(isect_ids, gaussian_ids, 0, 64)
    tile_bins = get_tile_bin_edges(num_intersects, isect_ids_sorted)
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins
