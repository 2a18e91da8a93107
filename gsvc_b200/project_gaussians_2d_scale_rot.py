"""project_gaussians_2d_scale_rot with the reference's signature
(gsplat/gsplat/project_gaussians_2d_scale_rot.py:12-151)."""
from __future__ import annotations

from typing import Tuple

from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
from .options import options


def project_gaussians_2d_scale_rot(means2d: Tensor, scales2d: Tensor, rotation: Tensor, img_height: int,
                                   img_width: int, tile_bounds: Tuple[int, int, int], clip_thresh: float = 0.01,
                                   ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Project 2D Gaussians given by scales (sx, sy) and a rotation angle.  Same outputs as
    project_gaussians_2d; differentiable w.r.t. means2d, scales2d and rotation."""
    out = _ProjectGaussians2dScaleRot.apply(means2d.contiguous(), scales2d.contiguous(), rotation.contiguous(),
                                            img_height, img_width, tile_bounds, clip_thresh)
    out[1]._gs2d_zero_depth = out[1]._version   # void once somebody writes into it
    return out


class _ProjectGaussians2dScaleRot(Function):
    @staticmethod
    def forward(ctx, means2d, scales2d, rotation, img_height, img_width, tile_bounds, clip_thresh=0.01):
        num_points = means2d.shape[-2]
        if num_points < 1 or means2d.shape[-1] != 2:
            raise ValueError(f"Invalid shape for means2d: {means2d.shape}")
        xys, depths, radii, conics, num_tiles_hit = _C.project_gaussians_2d_scale_rot_forward(
            num_points, means2d, scales2d, rotation, img_height, img_width, tile_bounds, clip_thresh)
        ctx.img_height = img_height
        ctx.img_width = img_width
        ctx.num_points = num_points
        ctx.analytic = bool(options.analytic_projection_grad)   # SURVEY 8f.4; default: reference formula
        ctx.save_for_backward(means2d, scales2d, rotation, radii, conics)
        ctx.mark_non_differentiable(radii, num_tiles_hit)
        return xys, depths, radii, conics, num_tiles_hit

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_num_tiles_hit):
        means2d, scales2d, rotation, radii, conics = ctx.saved_tensors
        _v_cov2d, v_mean2d, v_scale, v_rot = _C.project_gaussians_2d_scale_rot_backward(
            ctx.num_points, means2d, scales2d, rotation, ctx.img_height, ctx.img_width, radii, conics, v_xys,
            v_depths, v_conics, analytic=ctx.analytic)
        return v_mean2d, v_scale, v_rot, None, None, None, None
