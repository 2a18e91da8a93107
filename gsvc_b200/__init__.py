"""gsvc_b200 -- the gsplat 2D-Gaussian hot path of ac-freeman/GSVC, rebuilt for NVIDIA B200.

Public surface = the part of the reference's `gsplat` package that lies on the path
(gsplat/gsplat/__init__.py:20-47 of the reference): the three operators and the five binning
utilities, same names, same signatures.  `gsvc_b200.install_as_gsplat()` (or putting
gsvc_b200/dropin on PYTHONPATH) makes `import gsplat` resolve here, so GSVC's
GaussianSplats_Represent.py / GaussianSplats_Compress.py run unchanged.

Everything executes in libgs2d.so (hand-written sm_100a CUDA behind a C ABI, include/gs2d.h);
there is no CPU, Triton or PyTorch fallback -- without the library or a B200 every call raises.
"""
from __future__ import annotations

import sys
import warnings
from typing import Any

import torch

from .options import options
from .project_gaussians_2d import project_gaussians_2d
from .project_gaussians_2d_scale_rot import project_gaussians_2d_scale_rot
from .rasterize_sum import rasterize_gaussians_sum
from .utils import (
    bin_and_sort_gaussians,
    compute_cov2d_bounds,
    compute_cumulative_intersects,
    get_tile_bin_edges,
    map_gaussian_to_intersects,
)

__version__ = "0.1.3+b200"   # the reference vendors gsplat 0.1.3 (gsplat/gsplat/version.py:1)

__all__ = [
    "__version__",
    "project_gaussians_2d",
    "project_gaussians_2d_scale_rot",
    "rasterize_gaussians_sum",
    "bin_and_sort_gaussians",
    "compute_cumulative_intersects",
    "compute_cov2d_bounds",
    "get_tile_bin_edges",
    "map_gaussian_to_intersects",
    "ProjectGaussians2d",
    "ProjectGaussians2dScaleRot",
    "RasterizeGaussiansSum",
    "BinAndSortGaussians",
    "ComputeCumulativeIntersects",
    "ComputeCov2dBounds",
    "GetTileBinEdges",
    "MapGaussiansToIntersects",
    "install_as_gsplat",
    "options",
]


def _deprecated(name: str, new_name: str, fn):
    """the reference keeps Function-style aliases that warn and forward
    (gsplat/gsplat/__init__.py:50-212); same behaviour here"""

    class _Alias(torch.autograd.Function):
        @staticmethod
        def forward(ctx, *args, **kwargs):
            warnings.warn(f"{name} is deprecated, use {new_name} instead", DeprecationWarning)
            return fn(*args, **kwargs)

        @staticmethod
        def backward(ctx: Any, *grad_outputs: Any) -> Any:
            raise NotImplementedError

    _Alias.__name__ = name
    return _Alias


MapGaussiansToIntersects = _deprecated("MapGaussiansToIntersects", "map_gaussian_to_intersects", map_gaussian_to_intersects)
ComputeCumulativeIntersects = _deprecated("ComputeCumulativeIntersects", "compute_cumulative_intersects", compute_cumulative_intersects)
ComputeCov2dBounds = _deprecated("ComputeCov2dBounds", "compute_cov2d_bounds", compute_cov2d_bounds)
GetTileBinEdges = _deprecated("GetTileBinEdges", "get_tile_bin_edges", get_tile_bin_edges)
BinAndSortGaussians = _deprecated("BinAndSortGaussians", "bin_and_sort_gaussians", bin_and_sort_gaussians)
ProjectGaussians2d = _deprecated("ProjectGaussians2d", "project_gaussians_2d", project_gaussians_2d)
ProjectGaussians2dScaleRot = _deprecated("ProjectGaussians2dScaleRot", "project_gaussians_2d_scale_rot", project_gaussians_2d_scale_rot)
RasterizeGaussiansSum = _deprecated("RasterizeGaussiansSum", "rasterize_gaussians_sum", rasterize_gaussians_sum)


def install_as_gsplat() -> None:
    """Register this package under the name `gsplat` (and its sub-modules under the reference's
    module names) so `from gsplat.project_gaussians_2d import project_gaussians_2d` and
    `from gsplat.rasterize_sum import rasterize_gaussians_sum` (GaussianSplats_Represent.py:1-2,
    GaussianSplats_Compress.py:1-2) import this implementation."""
    me = sys.modules[__name__]
    sys.modules["gsplat"] = me
    for sub in ("project_gaussians_2d", "project_gaussians_2d_scale_rot", "rasterize_sum", "utils", "cuda"):
        sys.modules[f"gsplat.{sub}"] = sys.modules[f"{__name__}.{sub}"]
