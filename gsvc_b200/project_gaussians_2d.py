"""project_gaussians_2d with the reference's signature (gsplat/gsplat/project_gaussians_2d.py:12-141)."""
from __future__ import annotations

from typing import Tuple

from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
from .options import options


def project_gaussians_2d(means2d: Tensor, L_elements: Tensor, img_height: int, img_width: int,
                         tile_bounds: Tuple[int, int, int], clip_thresh: float = 0.01,
                         ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Project 2D Gaussians given by their Cholesky factor (l11, l21, l22).

    Returns (xys [N,2] pixels, depths [N] (all zero), radii [N] int32, conics [N,3],
    num_tiles_hit [N] int32); differentiable w.r.t. means2d and L_elements."""
    out = _ProjectGaussians2d.apply(means2d.contiguous(), L_elements.contiguous(), img_height, img_width,
                                    tile_bounds, clip_thresh)
    # depths are identically zero here (foward2d.cu:66-67): lets the rasteriser sort on tile id only
    out[1]._gs2d_zero_depth = out[1]._version   # void once somebody writes into it
    return out


class _ProjectGaussians2d(Function):
    @staticmethod
    def forward(ctx, means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh=0.01):
        num_points = means2d.shape[-2]
        xys, depths, radii, conics, num_tiles_hit = _C.project_gaussians_2d_forward(
            num_points, means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh)
        ctx.img_height = img_height
        ctx.img_width = img_width
        ctx.num_points = num_points
        ctx.analytic = bool(options.analytic_projection_grad)   # SURVEY 8f.4; default: reference formula
        ctx.save_for_backward(means2d, L_elements, radii, conics)
        ctx.mark_non_differentiable(radii, num_tiles_hit)
        return xys, depths, radii, conics, num_tiles_hit

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_num_tiles_hit):
        means2d, L_elements, radii, conics = ctx.saved_tensors
        _v_cov2d, v_mean2d, v_L_elements = _C.project_gaussians_2d_backward(
            ctx.num_points, means2d, L_elements, ctx.img_height, ctx.img_width, radii, conics, v_xys, v_depths,
            v_conics, analytic=ctx.analytic)
        return v_mean2d, v_L_elements, None, None, None, None
