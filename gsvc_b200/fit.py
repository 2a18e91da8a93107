"""Host-side mirror of the reference's per-frame fit loop (GaussianVideo_frame of
GaussianSplats_Represent.py:11-221 in the reference: parameters, activations, forward,
train_iter), reduced to what drives the hot path.  It exists so that bench.py and the tests can
run the loop BASELINE.json's config 2 names without the reference tree (which is not present on
the GPU box); GSVC's own model files use the package through `import gsplat` unchanged.

    FrameFit(H, W, n, device)            parameters initialised like the reference
    .forward()                           -> render [1,3,H,W] (clamped, NCHW)
    .train_iter(gt)                      -> (loss tensor, psnr float)   L2 + Adan + StepLR
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .adan import Adan
from .project_gaussians_2d import project_gaussians_2d
from .rasterize_sum import rasterize_gaussians_sum


class FrameFit(torch.nn.Module):
    def __init__(self, H: int, W: int, num_points: int, device, lr: float = 1e-3, seed: int = 1,
                 init: dict | None = None):
        super().__init__()
        self.H, self.W = int(H), int(W)
        self.tile_bounds = ((self.W + 15) // 16, (self.H + 15) // 16, 1)
        self.device = torch.device(device)
        if init is None:
            from . import synth
            init = synth.gaussians(num_points, seed=seed)
        # GaussianSplats_Represent.py:28-45: atanh-uniform means, uniform Cholesky, uniform colours
        self._xyz = torch.nn.Parameter(init["_xyz"].clone().to(self.device))
        self._cholesky = torch.nn.Parameter(init["_cholesky"].clone().to(self.device))
        self._features_dc = torch.nn.Parameter(init["_features_dc"].clone().to(self.device))
        self.register_buffer("cholesky_bound", torch.tensor([0.5, 0.0, 0.5], device=self.device).view(1, 3))
        self.register_buffer("background", torch.ones(3, device=self.device))
        self.register_buffer("opacity", torch.ones(num_points, 1, device=self.device))
        self.optimizer = Adan(self.parameters(), lr=lr)
        self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=20000, gamma=0.5)

    path_name = "autograd drop-in ops (project_gaussians_2d + rasterize_gaussians_sum) + torch glue + foreach Adan"

    @property
    def last_num_intersects(self) -> int:
        from .rasterize_sum import _device_ctx
        return _device_ctx(self.device).last_num_intersects

    def forward(self):
        xys, depths, radii, conics, nth = project_gaussians_2d(
            torch.tanh(self._xyz), self._cholesky + self.cholesky_bound, self.H, self.W, self.tile_bounds)
        out = rasterize_gaussians_sum(xys, depths, radii, conics, nth, self._features_dc, self.opacity, self.H,
                                      self.W, 16, 16, background=self.background, return_alpha=False)
        out = torch.clamp(out, 0, 1)
        return out.view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()

    def train_iter(self, gt_image: torch.Tensor):
        """one iteration of GaussianSplats_Represent.py:191-207 (no prune / add)"""
        image = self.forward()
        loss = F.mse_loss(image, gt_image)
        loss.backward()
        with torch.no_grad():
            psnr = 10 * math.log10(1.0 / loss.item())
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)
        self.scheduler.step()
        return loss, psnr
