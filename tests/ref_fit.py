"""The reference's fit iteration on the REFERENCE's own CUDA kernels (oracle/_ref, test
infrastructure): autograd Functions shaped like the reference's Python wrappers
(gsplat/gsplat/project_gaussians_2d.py:59-141, rasterize_sum.py:89-254 of the reference) around
tests/ref_gpu.py, driven by the same loop as gsvc_b200.fit.FrameFit.  Used by
`bench.py --impl reference-gpu` and by the trajectory parity test."""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

import ref_gpu
from gsvc_b200.adan import Adan


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


class RefFrameFit(torch.nn.Module):
    path_name = "reference CUDA kernels (oracle/_ref, sm_100a, --use_fast_math) + torch cumsum/.item()/sort/gather glue"
    last_num_intersects = 0

    def __init__(self, H, W, num_points, device, lr=1e-3, seed=1, init=None):
        super().__init__()
        from gsvc_b200 import synth
        self.H, self.W = H, W
        self.tb = ((W + 15) // 16, (H + 15) // 16, 1)
        self.device = torch.device(device)
        init = init or synth.gaussians(num_points, seed=seed)
        self._xyz = torch.nn.Parameter(init["_xyz"].clone().to(device))
        self._cholesky = torch.nn.Parameter(init["_cholesky"].clone().to(device))
        self._features_dc = torch.nn.Parameter(init["_features_dc"].clone().to(device))
        self.register_buffer("cholesky_bound", torch.tensor([0.5, 0.0, 0.5], device=device).view(1, 3))
        self.optimizer = Adan(self.parameters(), lr=lr)
        self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=20000, gamma=0.5)

    def forward(self):
        opacity = torch.ones(self._xyz.shape[0], 1).to(self.device)   # as the reference: new tensor every call
        xys, depths, radii, conics, nth = _RefProject.apply(torch.tanh(self._xyz), self._cholesky + self.cholesky_bound,
                                                            self.H, self.W, self.tb)
        out = _RefRasterSum.apply(xys, depths, radii, conics, nth, self._features_dc, opacity, self.H, self.W, self.tb)
        out = torch.clamp(out, 0, 1)
        return out.view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()

    def train_iter(self, gt_image):
        image = self.forward()
        loss = F.mse_loss(image, gt_image)
        loss.backward()
        with torch.no_grad():
            psnr = 10 * math.log10(1.0 / loss.item())
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)
        self.scheduler.step()
        return loss, psnr
