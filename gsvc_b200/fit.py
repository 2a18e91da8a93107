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


class FusedFrameFit(torch.nn.Module):
    """The same fit loop on the fused C-ABI entry point gs2d_fit_step (include/gs2d.h): four
    kernels per iteration, nothing synchronises the host until a loss is read.  Parameters keep
    the reference's names and shapes (`_xyz`, `_cholesky`, `_features_dc`), so a state dict moves
    freely between this class, FrameFit and the reference's GaussianVideo_frame.

        .train_iter(gt)          one iteration; returns (loss, psnr) -- reads the loss back
        .train_iters(gt, k)      k iterations, nothing read back; .loss(step) reads lazily
        .forward()               render [1,3,H,W] (clamped), sync-free
    gt: fp32 [1,3,H,W] / [3,H,W] in [0,1] (the reference loader's format) or uint8 [H,W,3]."""

    path_name = "fused C-ABI fit step (count/scatter binning, tile kernel fwd+L2+bwd, projection-bwd+Adan): 4 kernels/iter"

    def __init__(self, H: int, W: int, num_points: int, device, lr: float = 1e-3, seed: int = 1,
                 init: dict | None = None, betas=(0.98, 0.92, 0.99), eps: float = 1e-8, weight_decay: float = 0.0,
                 cap: int | None = None, lr_step_size: int = 20000, lr_gamma: float = 0.5):
        super().__init__()
        from . import _lib
        from . import cuda as _C
        self._lib, self._C = _lib, _C
        self.H, self.W, self.n = int(H), int(W), int(num_points)
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if init is None:
            from . import synth
            init = synth.gaussians(num_points, seed=seed)
        dev = self.device
        self._xyz = torch.nn.Parameter(init["_xyz"].clone().float().contiguous().to(dev), requires_grad=False)
        self._cholesky = torch.nn.Parameter(init["_cholesky"].clone().float().contiguous().to(dev), requires_grad=False)
        self._features_dc = torch.nn.Parameter(init["_features_dc"].clone().float().contiguous().to(dev),
                                               requires_grad=False)
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), tuple(betas), float(eps), float(weight_decay)
        self.lr_step_size, self.lr_gamma = int(lr_step_size), float(lr_gamma)
        self.step_index = 0
        self.cap = int(cap) if cap else max(8 * self.n, 1 << 18)
        self._alloc()

    def _alloc(self):
        lay = self._lib.FitLayout()
        lib = self._lib.load()
        self._lib.check(lib.gs2d_fit_layout_for(self.n, self.cap, self.H, self.W, lay), "fit_layout_for")
        self.layout = lay
        dev = self.device
        self.work = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=dev)
        if not hasattr(self, "adan_state"):
            self.adan_state = torch.empty((4, self.n, 8), dtype=torch.float32, device=dev)
            adan_ptr = self.adan_state.data_ptr()
        else:
            adan_ptr = None   # regrowing the workspace keeps the optimizer state
        with self._C._guard(dev):
            self._lib.check(lib.gs2d_fit_init(self.n, self.cap, self.H, self.W, self.work.data_ptr(), lay.total_bytes,
                                              adan_ptr, torch.cuda.current_stream(dev).cuda_stream), "fit_init")
        self._status = self.work[lay.off_status:lay.off_status + 64].view(torch.int32)
        self._loss = self.work[lay.off_loss:lay.off_loss + 8 * self._lib.FIT_LOSS_SLOTS].view(torch.float64)
        self._render_img = None
        self._counted_for = None   # nothing pre-counted in a fresh workspace

    def _target(self, gt: torch.Tensor):
        if gt.dtype == torch.uint8:
            if gt.dim() != 3 or gt.shape != (self.H, self.W, 3):
                raise ValueError("uint8 target must be [H,W,3]")
            return gt.contiguous(), self._lib.TARGET_U8_HWC
        if gt.dtype != torch.float32 or gt.numel() != 3 * self.H * self.W:
            raise ValueError("target must be fp32 [1,3,H,W] / [3,H,W] or uint8 [H,W,3]")
        return gt.contiguous(), self._lib.TARGET_F32_CHW

    def current_lr(self) -> float:
        # torch.optim.lr_scheduler.StepLR(step_size=20000, gamma=0.5), GaussianSplats_Represent.py:52
        return self.lr * self.lr_gamma ** (self.step_index // self.lr_step_size)

    def _issue(self, gt, fmt, out_img=None, grads_out=None, ids_out=None):
        self.step_index += 1
        b1, b2, b3 = self.betas
        lib = self._lib.load()
        self._lib.check(lib.gs2d_fit_step(
            self.n, self.cap, self.H, self.W, self._xyz.data_ptr(), self._cholesky.data_ptr(),
            self._features_dc.data_ptr(), self.adan_state.data_ptr(), gt.data_ptr(), fmt, self.work.data_ptr(),
            self.layout.total_bytes, self.step_index, self.current_lr(), b1, b2, b3, self.eps, self.weight_decay,
            1 if self._precount_valid() else 0, None if out_img is None else out_img.data_ptr(), None if grads_out is None else grads_out.data_ptr(),
            None if ids_out is None else ids_out.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream),
            "fit_step")

    def _versions(self):
        return (self._xyz._version, self._cholesky._version, self._features_dc._version,
                self._xyz.data_ptr(), self._cholesky.data_ptr(), self._features_dc.data_ptr())

    def _precount_valid(self) -> bool:
        """the previous step pre-counted the tiles of the parameters it wrote; that is still valid
        unless someone modified the parameter tensors since (torch bumps ._version on in-place ops)"""
        ok = self._counted_for == self._versions()
        return ok

    def train_iters(self, gt_image: torch.Tensor, iters: int = 1, **dbg):
        gt, fmt = self._target(gt_image)
        if gt.device != self.device:
            raise RuntimeError("target frame must live on the model's device")
        with self._C._guard(self.device):
            for _ in range(int(iters)):
                self._issue(gt, fmt, **dbg)
                self._counted_for = self._versions()

    def status(self):
        """(num_intersects, overflow_step, max_per_tile) of the last issued step; synchronises"""
        s = self._status.cpu()
        return int(s[0]), int(s[1]), int(s[3])

    def loss(self, step: int | None = None) -> float:
        """mean squared error of `step` (default: the last one); synchronises.  A step's slot
        is recycled FIT_LOSS_SLOTS - 1 steps later."""
        step = self.step_index if step is None else int(step)
        if not (self.step_index - (self._lib.FIT_LOSS_SLOTS - 2) <= step <= self.step_index):
            raise ValueError("that step's loss slot has been recycled")
        both = torch.cat([self._loss[step % self._lib.FIT_LOSS_SLOTS].view(1),
                          self._status[1].to(torch.float64).view(1)]).cpu()
        if both[1] != 0:
            self._recover(int(both[1]))
            raise RuntimeError(
                f"intersections exceeded the workspace capacity at step {int(both[1])}; the workspace was regrown "
                f"to {self.cap} and the model rewound to step {self.step_index}: re-issue the remaining iterations")
        return float(both[0]) / (3.0 * self.H * self.W)

    def _recover(self, overflow_step: int):
        # steps >= overflow_step were no-ops (parameters and optimizer state untouched)
        self.step_index = overflow_step - 1
        self.cap *= 2
        self._alloc()

    def train_iter(self, gt_image: torch.Tensor):
        self.train_iters(gt_image, 1)
        loss = self.loss()
        return loss, 10 * math.log10(1.0 / loss)

    @property
    def last_num_intersects(self) -> int:
        return self.status()[0]

    @torch.no_grad()
    def forward(self):
        if self._render_img is None:
            lay = self._lib.FitLayout()
            lib = self._lib.load()
            self._lib.check(lib.gs2d_fit_layout_for(self.n, self.cap, self.H, self.W, lay), "fit_layout_for")
            self._render_work = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=self.device)
            with self._C._guard(self.device):
                self._lib.check(lib.gs2d_fit_init(self.n, self.cap, self.H, self.W, self._render_work.data_ptr(),
                                                  lay.total_bytes, None,
                                                  torch.cuda.current_stream(self.device).cuda_stream), "fit_init")
            self._render_img = True
        out = torch.empty((1, 3, self.H, self.W), dtype=torch.float32, device=self.device)
        with self._C._guard(self.device):
            self._lib.check(self._lib.load().gs2d_render_chol(
                self.n, self.cap, self.H, self.W, self._xyz.data_ptr(), self._cholesky.data_ptr(),
                self._features_dc.data_ptr(), None, 1, self._render_work.data_ptr(), self._render_work.numel(),
                out.data_ptr(), 1, 1, None, None, torch.cuda.current_stream(self.device).cuda_stream), "render_chol")
        return out
