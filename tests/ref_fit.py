"""The reference's fit iteration on the REFERENCE's own CUDA kernels (oracle/_ref, test
infrastructure): autograd Functions shaped like the reference's Python wrappers
(gsplat/gsplat/project_gaussians_2d.py:59-141, rasterize_sum.py:89-254 of the reference) around
tests/ref_gpu.py, driven by the same loop as gsvc_b200.fit.FrameFit.  Used by
`bench.py --impl reference-gpu` and by the trajectory parity test."""
from __future__ import annotations

import torch

try:
    from tests import ref_gpu
except ImportError:          # bench.py puts tests/ itself on sys.path
    import ref_gpu
from gsvc_b200.fit import FrameFit


class _RefProject(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, L, H, W, tb):
        xys, depths, radii, conics, nth = ref_gpu.project_gaussians_2d_forward(means2d, L, H, W, tb)
        ctx.H, ctx.W = H, W
        ctx.save_for_backward(means2d, L, radii, conics)
        ctx.mark_non_differentiable(radii, nth)
        return xys, depths, radii, conics, nth

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_nth):
        means2d, L, radii, conics = ctx.saved_tensors
        _, v_mean, v_L = ref_gpu.project_gaussians_2d_backward(means2d, L, ctx.H, ctx.W, radii, conics,
                                                               v_xys.contiguous(), v_conics.contiguous())
        return v_mean, v_L, None, None, None


class _RefRasterSum(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xys, depths, radii, conics, nth, colors, opacity, H, W, tb):
        n = xys.size(0)
        b = ref_gpu.bin_and_sort(n, xys, depths, radii, nth, tb, stable=False)   # .item() sync + torch.sort
        ctx.H, ctx.W, ctx.I = H, W, b["num_intersects"]
        if ctx.I < 1:
            ctx.save_for_backward(xys, conics, colors, opacity)
            return torch.ones(H, W, 3, device=xys.device)
        out, Ts, fidx = ref_gpu.rasterize_sum_forward(H, W, b["gaussian_ids_sorted"], b["tile_bins"], xys, conics,
                                                      colors, opacity)
        ctx.save_for_backward(xys, conics, colors, opacity, b["gaussian_ids_sorted"], b["tile_bins"], Ts, fidx)
        return out

    @staticmethod
    def backward(ctx, v_out):
        s = ctx.saved_tensors
        if ctx.I < 1:
            return (torch.zeros_like(s[0]), None, None, torch.zeros_like(s[1]), None, torch.zeros_like(s[2]),
                    torch.zeros_like(s[3]), None, None, None)
        xys, conics, colors, opacity, ids, bins, Ts, fidx = s
        v_xy, v_conic, v_col, v_op = ref_gpu.rasterize_sum_backward(ctx.H, ctx.W, ids, bins, xys, conics, colors,
                                                                    opacity, Ts, fidx, v_out)
        return v_xy, None, None, v_conic, None, v_col, v_op, None, None, None


class RefFrameFit(FrameFit):
    """gsvc_b200.fit.FrameFit (the loop, rgb_W and the prune / add control, pinned against the
    reference's model class by tests/test_density_cpu.py) with the two operators swapped for the
    reference's own kernels"""

    path_name = "reference CUDA kernels (oracle/_ref, sm_100a, --use_fast_math) + torch cumsum/.item()/sort/gather glue"
    last_num_intersects = 0

    def _render(self, means2d, L_elements, colors):
        opacity = torch.ones(means2d.shape[0], 1).to(self.device)   # as the reference: new tensor every call
        xys, depths, radii, conics, nth = _RefProject.apply(means2d, L_elements, self.H, self.W, self.tile_bounds)
        return _RefRasterSum.apply(xys, depths, radii, conics, nth, colors, opacity, self.H, self.W, self.tile_bounds)
