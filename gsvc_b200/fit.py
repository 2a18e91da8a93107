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
import os
from typing import Optional

import torch
import torch.nn.functional as F

from .adan import Adan
from .project_gaussians_2d import project_gaussians_2d
from .rasterize_sum import rasterize_gaussians_sum


class FrameFit(torch.nn.Module):
    """GaussianVideo_frame on the drop-in operators.  With isdensity / isremoval the colour weight
    rgb_W is learnable and the prune / add control of GaussianSplats_Represent.py:98-172 runs
    between backward() and the optimizer step, as train_iter :199-204 does (schedule:
    gsvc_b200.density.DensityPlan)."""

    def __init__(self, H: int, W: int, num_points: int, device, lr: float = 1e-3, seed: int = 1,
                 init: dict | None = None, isdensity: bool = False, isremoval: bool = False,
                 removal_rate: float = 0.1, max_num_points: int | None = None, densification_interval: int = 100,
                 plan=None):
        super().__init__()
        from .density import DensityPlan
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
        n = self._xyz.shape[0]
        self.plan = plan or DensityPlan(max_num_points or n, removal_rate, densification_interval, isdensity, isremoval)
        if self.plan.learns_rgb_w:      # :31-36
            w0 = init["rgb_W"].clone().float() if "rgb_W" in init else torch.full((n, 1), self.plan.rgb_w_init)
            self.rgb_W = torch.nn.Parameter(w0.to(self.device))
        else:
            self.register_buffer("rgb_W", torch.ones((n, 1), device=self.device))
        self.register_buffer("cholesky_bound", torch.tensor([0.5, 0.0, 0.5], device=self.device).view(1, 3))
        self.register_buffer("background", torch.ones(3, device=self.device))
        self.lr = lr
        self.optimizer = Adan(self.parameters(), lr=lr)
        self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=20000, gamma=0.5)
        self.iter = 0

    path_name = "autograd drop-in ops (project_gaussians_2d + rasterize_gaussians_sum) + torch glue + foreach Adan"

    @property
    def last_num_intersects(self) -> int:
        from .rasterize_sum import _device_ctx
        return _device_ctx(self.device).last_num_intersects

    @property
    def get_features(self):
        return self._features_dc * self.rgb_W      # :62-63

    def _render(self, means2d, L_elements, colors):
        """[H,W,3] unclamped; the one place that touches the operators (tests swap it out)"""
        xys, depths, radii, conics, nth = project_gaussians_2d(means2d, L_elements, self.H, self.W, self.tile_bounds)
        opacity = torch.ones(means2d.shape[0], 1, device=means2d.device)
        return rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, opacity, self.H, self.W, 16, 16,
                                       background=self.background, return_alpha=False)

    def forward(self):
        out = self._render(torch.tanh(self._xyz), self._cholesky + self.cholesky_bound, self.get_features)
        out = torch.clamp(out, 0, 1)
        return out.view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()

    def frame_state(self) -> dict:
        """train_video_Represent.py:107-114: what a fitted frame hands to the next one"""
        return {"_xyz": self._xyz.detach().clone(), "_cholesky": self._cholesky.detach().clone(),
                "_features_dc": self.get_features.detach().clone()}

    def _reattach(self):
        for group in self.optimizer.param_groups:       # :115-116
            group["params"] = [p for p in self.parameters() if p.requires_grad]

    @torch.no_grad()
    def _prune(self, count: int):
        # torch.norm + torch.sort + boolean mask, :103-114.  torch.sort leaves the order of equal
        # keys open; this package fixes it to lowest index first (as gs2d_prune_lowest does)
        _, order = torch.sort(torch.norm(self.rgb_W, dim=1), stable=True)
        keep = torch.ones(self._xyz.shape[0], dtype=torch.bool, device=self._xyz.device)
        keep[order[:count]] = False
        self.last_keep = keep
        self._xyz = torch.nn.Parameter(self._xyz[keep])
        self._cholesky = torch.nn.Parameter(self._cholesky[keep])
        self._features_dc = torch.nn.Parameter(self._features_dc[keep])
        self.rgb_W = torch.nn.Parameter(self.rgb_W[keep])

    @torch.no_grad()
    def _add(self, count: int):
        # :134-144 -- drawn from the CPU generator in this order, then moved to the device
        dev = self._xyz.device
        new_xyz = torch.atanh(2 * (torch.rand(count, 2) - 0.5)).to(dev)
        new_cholesky = torch.rand(count, 3).to(dev)
        new_features_dc = torch.rand(count, 3).to(dev)
        new_rgb_W = 0.01 * torch.ones(count, 1).to(dev)
        self._xyz = torch.nn.Parameter(torch.cat((self._xyz, new_xyz), dim=0))
        self._cholesky = torch.nn.Parameter(torch.cat((self._cholesky, new_cholesky), dim=0))
        self._features_dc = torch.nn.Parameter(torch.cat((self._features_dc, new_features_dc), dim=0))
        self.rgb_W = torch.nn.Parameter(torch.cat((self.rgb_W, new_rgb_W), dim=0))

    def _control(self, it: int):
        act = self.plan.action(it, self._xyz.shape[0])
        if act.kind == "prune":
            self._prune(act.count)
        elif act.kind == "add":
            self._add(act.count)
        if act.rebuild:
            self.optimizer = Adan(self.parameters(), lr=self.lr)       # update_optimizer(), :93-97
        elif act.replaces_parameters:
            self._reattach()
        return act

    def train_iter(self, gt_image: torch.Tensor, iter: int | None = None):
        """one iteration of GaussianSplats_Represent.py:191-207"""
        self.iter = self.iter + 1 if iter is None else int(iter)
        image = self.forward()
        loss = F.mse_loss(image, gt_image)
        loss.backward()
        with torch.no_grad():
            psnr = 10 * math.log10(1.0 / loss.item())
        if self.plan.learns_rgb_w:
            self._control(self.iter)
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)
        self.scheduler.step()
        return loss, psnr

    # ---- the same iteration as one CUDA-graph replay ----------------------------------------------------
    def train_iter_graphed(self, gt_image: torch.Tensor, read_loss: bool = True):
        """train_iter with the whole iteration -- both operators, the torch glue around them, their
        backwards, the Adan step -- captured ONCE in a CUDA graph and replayed: one host call per
        iteration instead of ~45 launches (the operators are sync-free and capturable; the reference's
        cannot be captured, they block on .item(), utils.py:117).  Same arithmetic as train_iter.
        The first call of a configuration runs its iteration eagerly (on a side stream, as torch's
        capture recipe asks) and captures; iterations in which the prune / add control acts, and the
        one after them, run eagerly too and the graph is captured again (new parameter tensors).
        read_loss=False skips the per-iteration loss.item() (psnr is then None)."""
        it = self.iter + 1
        act = self.plan.action(it, self._xyz.shape[0]) if self.plan.learns_rgb_w else _NO_ACTION
        lr = self.optimizer.param_groups[0]["lr"]
        key = (id(self.optimizer), tuple(p.data_ptr() for p in self.parameters()), lr, tuple(gt_image.shape))
        if act.kind != "none" or act.rebuild:
            self._graph = None
            return self.train_iter(gt_image)
        if getattr(self, "_graph", None) is None or self._graph_key != key:
            return self._capture_iteration(gt_image, key, read_loss)
        if gt_image.data_ptr() != self._g_gt.data_ptr():
            self._g_gt.copy_(gt_image, non_blocking=True)
        self._graph.replay()
        self.iter = it
        self.optimizer.param_groups[0]["step"] += 1        # the replay bumped the device-side counter
        self.scheduler.step()
        if not read_loss:
            return self._g_loss, None
        return self._g_loss, 10 * math.log10(1.0 / self._g_loss.item())

    @torch.no_grad()
    def forward_graphed(self):
        """forward() (the render of train_video_Represent.py:101-106) as one CUDA-graph replay; the
        returned image is the graph's static output, overwritten by the next replay"""
        key = tuple(p.data_ptr() for p in self.parameters())
        if getattr(self, "_fgraph", None) is None or self._fgraph_key != key:
            cur = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                self.forward()                              # sizes the operator's buffers outside the capture
            cur.wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = self.forward()
            self._fgraph, self._fgraph_key, self._fgraph_out = graph, key, out
        self._fgraph.replay()
        return self._fgraph_out

    def _capture_iteration(self, gt_image, key, read_loss):
        self.optimizer.capturable = True
        cur = torch.cuda.current_stream(self.device)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            loss, psnr = self.train_iter(gt_image)          # this call's iteration (reads the loss: drains the stream)
        cur.wait_stream(side)
        eager = (loss.detach(), psnr)
        del loss                                            # frees the autograd graph and with it the leaves' accumulators
        key = (key[0], tuple(p.data_ptr() for p in self.parameters()), self.optimizer.param_groups[0]["lr"], key[3])
        self._g_gt = gt_image.detach().clone()
        group = self.optimizer.param_groups[0]
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            image = self.forward()
            g_loss = F.mse_loss(image, self._g_gt)
            g_loss.backward()
            self.optimizer.step()
            self.optimizer.zero_grad(set_to_none=True)
        group["step"] -= 1                                   # the capture ran no step; neither counter moved
        group["step_dev"].fill_(group["step"])
        self._graph, self._graph_key, self._g_loss = graph, key, g_loss.detach()
        return eager


class GS2DOverflow(RuntimeError):
    """A fused step found its workspace too small (more intersections than `cap`, or -- on tile slots --
    a tile with more than 256 Gaussians): that step and every later one were no-ops.  By the time this is
    raised the model has already rewound itself to the last step that really happened and regrown /
    switched its workspace: re-issue the remaining iterations (FusedFrameFit.train_until does)."""


class _NoAction:
    kind, count, rebuild, replaces_parameters = "none", 0, False, False


_NO_ACTION = _NoAction()


class FusedFrameFit(torch.nn.Module):
    """The same fit loop on the fused C-ABI entry point gs2d_fit_step (include/gs2d.h): four
    kernels per iteration, nothing synchronises the host until a loss is read.  Parameters keep
    the reference's names and shapes (`_xyz`, `_cholesky`, `_features_dc`), so a state dict moves
    freely between this class, FrameFit and the reference's GaussianVideo_frame.

        .train_iter(gt)          one iteration; returns (loss, psnr) -- reads the loss back
        .train_iters(gt, k)      k iterations, nothing read back; .loss(step) reads lazily
        .forward()               render [1,3,H,W] (clamped), sync-free
    gt: fp32 [1,3,H,W] / [3,H,W] in [0,1] (the reference loader's format), uint8 RGB [H,W,3], or
    the raw uint8 I420 frame [H*3/2,W] of the .yuv file (decoded inside the tile kernel exactly as
    the reference's cv2 loader would)."""

    @property
    def path_name(self) -> str:
        if self.slots:
            return ("fused C-ABI fit step on tile slots: tile kernel (fwd + clamp/L2 + bwd) and update kernel "
                    "(projection-bwd + Adan + the next step's projection and records), PDL: 2 kernels/iter")
        return ("fused C-ABI fit step on tile ranges: scatter-pack, tile kernel (fwd + clamp/L2 + bwd), update kernel "
                "(projection-bwd + Adan + next count/scan): 3 kernels/iter")

    def __init__(self, H: int, W: int, num_points: int, device, lr: float = 1e-3, seed: int = 1,
                 init: dict | None = None, betas=(0.98, 0.92, 0.99), eps: float = 1e-8, weight_decay: float = 0.0,
                 cap: int | None = None, lr_step_size: int = 20000, lr_gamma: float = 0.5, isdensity: bool = False,
                 isremoval: bool = False, removal_rate: float = 0.1, max_num_points: int | None = None,
                 densification_interval: int = 100, plan=None, slots: bool = True, analytic_grad: bool | None = None,
                 max_per_tile: int | None = None):
        """slots: fixed 256-record slots per tile instead of tile ranges (include/gs2d.h,
        gs2d_fit_options.slots): two kernels per iteration instead of three.  The first tile that
        would need a 257th record switches the model to ranges for good (same results, see _recover).
        analytic_grad: the projection backward's analytic gradient instead of the reference's formula
        (SURVEY Q12 / 8f.4); None = gsvc_b200.options.analytic_projection_grad (default off).
        max_per_tile: 256 = the reference renders (and fits) only the first 256 Gaussians of a tile
        (SURVEY Q1); 0 = all of them, in batches of 256 (a tile that fills its 256 slots moves the model to
        tile ranges, where that applies); None = gsvc_b200.options.max_per_tile (default 256)."""
        super().__init__()
        from .options import options as _options
        self.analytic_grad = bool(_options.analytic_projection_grad if analytic_grad is None else analytic_grad)
        mpt = int(_options.max_per_tile if max_per_tile is None else max_per_tile)
        if mpt not in (0, 256):
            raise ValueError("max_per_tile must be 256 (the reference) or 0 (no limit)")
        self.all_per_tile = mpt == 0
        import ctypes
        from . import _lib
        from . import cuda as _C
        from .density import DensityPlan, OptimizerClock
        self._lib, self._C, self._ctypes = _lib, _C, ctypes
        self.H, self.W, self.n = int(H), int(W), int(num_points)
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if init is None:
            from . import synth
            init = synth.gaussians(num_points, seed=seed)
        dev = self.device
        self._pack(init["_xyz"].float().to(dev), init["_cholesky"].float().to(dev), init["_features_dc"].float().to(dev))
        # prune / add control (GaussianSplats_Represent.py:98-172): rgb_W is a parameter only then (:31-36)
        self.plan = plan or DensityPlan(max_num_points or self.n, removal_rate, densification_interval, isdensity,
                                        isremoval)
        if self.plan.learns_rgb_w:
            w0 = init["rgb_W"].clone().float() if "rgb_W" in init else torch.full((self.n, 1), self.plan.rgb_w_init)
            self.rgb_W = torch.nn.Parameter(w0.contiguous().to(dev), requires_grad=False)
        else:
            self.rgb_W = None
        self.clock = OptimizerClock()
        self.rng = None   # torch.Generator for the random add (None = torch's global CPU generator)
        self._opts = _lib.FitOptions()
        self._cdll = _lib.load()
        self._tile_rows = (self.H + 15) // 16
        self._both_phases = _lib.FIT_PHASE_GRADS | _lib.FIT_PHASE_UPDATE
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), tuple(betas), float(eps), float(weight_decay)
        self.lr_step_size, self.lr_gamma = int(lr_step_size), float(lr_gamma)
        self.step_index = 0
        self._ckpt = (0, 0, True, False)     # (iteration, clock.step, clock.fresh, clock.lr_frozen) at the last sync
        self.slots = bool(slots)
        self._cap_ranges = int(cap) if cap else max(8 * max(self.n, self.plan.max_num_points), 1 << 18)
        tiles = ((self.H + 15) // 16) * ((self.W + 15) // 16)
        self.cap = max(self._cap_ranges, tiles * 256) if self.slots else self._cap_ranges
        self._alloc()

    def _pack(self, xyz: torch.Tensor, chol: torch.Tensor, feat: torch.Tensor, into: torch.Tensor | None = None):
        """The master copy of the parameters is ONE fp32 [n][8] buffer (rows: xyz 2 | chol 3 | features 3;
        gs2d_fit_options.params8): every per-Gaussian kernel moves a Gaussian as two aligned 16-byte
        accesses.  `_xyz`, `_cholesky`, `_features_dc` keep the reference's names and shapes
        (train_video_Represent.py:107-114) as column views of it; state_dict() / frame_state() hand out
        contiguous copies."""
        n = int(xyz.shape[0])
        p8 = torch.empty((n, 8), dtype=torch.float32, device=self.device) if into is None else into
        p8[:, 0:2] = xyz
        p8[:, 2:5] = chol
        p8[:, 5:8] = feat
        self._params8 = p8
        self._xyz = torch.nn.Parameter(p8[:, 0:2], requires_grad=False)
        self._cholesky = torch.nn.Parameter(p8[:, 2:5], requires_grad=False)
        self._features_dc = torch.nn.Parameter(p8[:, 5:8], requires_grad=False)
        self.n = n

    def _param_ptrs(self):
        """(params8, xyz, chol, feat) device addresses; repacks first if somebody assigned a new tensor
        to one of the three attributes instead of writing into it"""
        base = self._params8.data_ptr()
        if (self._xyz.data_ptr() != base or self._cholesky.data_ptr() != base + 8
                or self._features_dc.data_ptr() != base + 20 or self._xyz.stride(0) != 8):
            with torch.no_grad():
                self._pack(self._xyz.detach().float().to(self.device), self._cholesky.detach().float().to(self.device),
                           self._features_dc.detach().float().to(self.device))
            self._counted_for = None
            base = self._params8.data_ptr()
        return base, base, base + 8, base + 20

    def state_dict(self, *args, **kwargs):
        sd = super().state_dict(*args, **kwargs)
        for k in list(sd):
            if k.split(".")[-1] in ("_xyz", "_cholesky", "_features_dc"):
                sd[k] = sd[k].contiguous()     # a column view would drag the whole [n][8] storage into torch.save
        return sd

    def _alloc(self):
        lay = self._lib.FitLayout()
        lib = self._lib.load()
        self._lib.check(lib.gs2d_fit_layout_for(self.n, self.cap, self.H, self.W, lay), "fit_layout_for")
        self.layout = lay
        dev = self.device
        self.work = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=dev)
        if getattr(self, "adan_state", None) is None or self.adan_state.shape[1] != self.n:
            # new model or new Gaussian count: fresh optimizer state (regrowing the workspace keeps it)
            self.adan_state = torch.zeros((4, self.n, 8), dtype=torch.float32, device=dev)
            self.adan_state_w = torch.zeros((4, self.n), dtype=torch.float32, device=dev) if self.rgb_W is not None else None
        with self._C._guard(dev):
            self._lib.check(lib.gs2d_fit_init(self.n, self.cap, self.H, self.W, self.work.data_ptr(), lay.total_bytes,
                                              None, torch.cuda.current_stream(dev).cuda_stream), "fit_init")
        self._status = self.work[lay.off_status:lay.off_status + 64].view(torch.int32)
        self._loss = self.work[lay.off_loss:lay.off_loss + 8 * self._lib.FIT_LOSS_SLOTS].view(torch.float64)
        if not hasattr(self, "_report"):
            # ring of gs2d_fit_report records (24 B each) in pinned host memory, written by the device
            self._report = torch.zeros((self._lib.FIT_LOSS_SLOTS, 6), dtype=torch.int32).pin_memory()
            self._report[:, 5] = -1
            self._report_np = self._report.numpy()
            self._report_f64 = self._report_np.view("float64")     # [slots, 3]; column 0 = loss_sum
            self._report_ptr = self._report.data_ptr()
        self._render_img = None
        self._counted_for = None   # nothing pre-counted in a fresh workspace

    def _target(self, gt: torch.Tensor):
        if gt.dtype == torch.uint8:
            if gt.numel() * 2 == 3 * self.H * self.W and gt.dim() <= 2:
                # the raw I420 frame of the .yuv file ([H*3/2, W] as utils.py:148 shapes it, or flat)
                if self.H % 2 or self.W % 2:
                    raise ValueError("an I420 target needs even width and height")
                return gt.contiguous(), self._lib.TARGET_I420
            if gt.dim() != 3 or gt.shape != (self.H, self.W, 3):
                raise ValueError("uint8 target must be RGB [H,W,3] or a raw I420 frame [H*3/2,W]")
            return gt.contiguous(), self._lib.TARGET_U8_HWC
        if gt.dtype != torch.float32 or gt.numel() != 3 * self.H * self.W:
            raise ValueError("target must be fp32 [1,3,H,W] / [3,H,W] or uint8 [H,W,3]")
        return gt.contiguous(), self._lib.TARGET_F32_CHW

    def current_lr(self) -> float:
        # torch.optim.lr_scheduler.StepLR(step_size=20000, gamma=0.5), GaussianSplats_Represent.py:52;
        # the optimizer update_optimizer() builds (:93-97) has no scheduler: constant lr from then on
        # iteration `it` (1-based, = step_index while it is being issued) runs optimizer.step() BEFORE the
        # it-th scheduler.step() (GaussianSplats_Represent.py:203-206), i.e. at lr * gamma^((it-1) // step)
        if self.clock.lr_frozen:
            return self.lr
        return self.lr * self.lr_gamma ** (max(self.step_index - 1, 0) // self.lr_step_size)

    def _options(self, adan_step: int, fresh: bool, skip: bool, w_grad_out=None):
        o = self._opts
        if self.rgb_W is not None:
            o.rgb_w, o.adan_state_w = self.rgb_W.data_ptr(), self.adan_state_w.data_ptr()
            o.rgb_w_grad_out = None if w_grad_out is None else w_grad_out.data_ptr()
        o.adan_step, o.adan_fresh, o.skip_update = int(adan_step), int(fresh), int(skip)
        o.slots = int(self.slots)
        o.analytic_grad = int(self.analytic_grad)
        o.all_per_tile = int(self.all_per_tile and not self.slots)   # slots hold 256: a fuller tile moves the model to ranges
        o.params8 = self._params8.data_ptr()
        return o

    def _issue(self, gt, fmt, out_img=None, grads_out=None, ids_out=None, w_grad_out=None, stream=None):
        """gt: the target on the device (tensor, or a raw device address)"""
        self.step_index += 1
        act = self.plan.action(self.step_index, self.n) if self.rgb_W is not None else _NO_ACTION
        if act.rebuild or act.replaces_parameters:
            self._control_sync()                    # the one host wait of the control: once per interval
        if act.rebuild and not act.replaces_parameters:
            self._zero_optimizer_state()            # update_optimizer() precedes this iteration's step
        adan_step, fresh, skip = self.clock.begin(act)
        _p8, p_xyz, p_chol, p_feat = self._param_ptrs()
        opts = self._options(adan_step, fresh, skip, w_grad_out)
        b1, b2, b3 = self.betas
        slot = self.step_index % self._lib.FIT_LOSS_SLOTS
        rc = self._cdll.gs2d_fit_step_band(
            self.n, self.cap, self.H, self.W, p_xyz, p_chol,
            p_feat, self.adan_state.data_ptr(), gt if isinstance(gt, int) else gt.data_ptr(),
            fmt, self.work.data_ptr(),
            self.layout.total_bytes, self.step_index, self.current_lr(), b1, b2, b3, self.eps, self.weight_decay,
            1 if self._precount_valid() else 0, 0, self._tile_rows,
            self._both_phases, None, self._ctypes.addressof(opts),
            None if out_img is None else out_img.data_ptr(),
            None if grads_out is None else grads_out.data_ptr(), None if ids_out is None else ids_out.data_ptr(),
            self._report_ptr + 24 * slot,
            torch.cuda.current_stream(self.device).cuda_stream if stream is None else stream)
        if rc:
            self._lib.check(rc, "fit_step")
        if skip:
            self._counted_for = None                # a skipped update pre-counts nothing
            self._replace_parameters(act)
        else:
            self._counted_for = self._versions()
        if act.rebuild or act.replaces_parameters:
            self._ckpt = (self.step_index, self.clock.step, self.clock.fresh, self.clock.lr_frozen)

    def _control_sync(self):
        """before a prune / add / optimizer rebuild: all earlier steps must have really happened
        (an overflowed workspace turns steps into no-ops; see _recover)"""
        if self.step_index <= 1:
            return
        torch.cuda.current_stream(self.device).synchronize()
        prev = self.step_index - 1
        _, _i, overflow, _m = self._read_report(prev)
        if overflow != 0:
            self.step_index = prev
            raise GS2DOverflow(self._recover(overflow))

    def _zero_optimizer_state(self):
        self.adan_state.zero_()
        if self.adan_state_w is not None:
            self.adan_state_w.zero_()

    @torch.no_grad()
    def _replace_parameters(self, act):
        """the prune (gs2d_prune_lowest) or the random add of GaussianSplats_Represent.py:134-144,
        after the step that only reported the loss; optimizer state starts afresh for every row"""
        dev, n = self.device, self.n
        if act.kind == "prune":
            m = n - act.count
            if m < 1:
                raise RuntimeError("the prune would remove every Gaussian")
            lib = self._lib.load()
            scratch = torch.empty((lib.gs2d_prune_scratch_bytes(n),), dtype=torch.uint8, device=dev)
            new = [torch.empty((m, c), dtype=torch.float32, device=dev) for c in (1, 2, 3, 3)]
            self.last_keep_index = torch.empty((m,), dtype=torch.int32, device=dev)
            src = [t.detach().contiguous() for t in (self._xyz, self._cholesky, self._features_dc)]   # [n,2] [n,3] [n,3]
            self._lib.check(lib.gs2d_prune_lowest(
                n, act.count, self.rgb_W.data_ptr(), src[0].data_ptr(), src[1].data_ptr(),
                src[2].data_ptr(), new[0].data_ptr(), new[1].data_ptr(), new[2].data_ptr(),
                new[3].data_ptr(), self.last_keep_index.data_ptr(), scratch.data_ptr(), scratch.numel(),
                torch.cuda.current_stream(dev).cuda_stream), "prune_lowest")
        else:
            # same draws, in the same order, from the global CPU generator as the reference
            c = act.count
            g = self.rng   # None: the global CPU generator, as the reference; gop.fit_gop hands every frame its own
            add = [torch.atanh(2 * (torch.rand(c, 2, generator=g) - 0.5)), torch.rand(c, 3, generator=g),
                   torch.rand(c, 3, generator=g), 0.01 * torch.ones(c, 1)]
            old = [self._xyz, self._cholesky, self._features_dc, self.rgb_W]
            cat = [torch.cat((o, a.to(dev)), dim=0) for o, a in zip(old, add)]
            new = [cat[3], cat[0], cat[1], cat[2]]
        self.rgb_W = torch.nn.Parameter(new[0], requires_grad=False)
        self._pack(new[1], new[2], new[3])
        self.adan_state = None                      # new tensors, new (zero) optimizer state
        self._alloc()

    def _versions(self):
        w = self.rgb_W
        return (self._xyz._version, self._cholesky._version, self._features_dc._version,
                self._xyz.data_ptr(), self._cholesky.data_ptr(), self._features_dc.data_ptr(),
                None if w is None else (w._version, w.data_ptr()))

    def _precount_valid(self) -> bool:
        """the previous step pre-counted the tiles of the parameters it wrote; that is still valid
        unless someone modified the parameter tensors since (torch bumps ._version on in-place ops)"""
        return self._counted_for == self._versions()

    #: convert a resident target frame once into the tile kernel's own order (gs2d_fit_target_prepare) instead
    #: of decoding its bytes in every iteration; False keeps the raw format (what fit_stream always does)
    prepare_targets = os.environ.get("GS2D_NO_PREPARE") != "1"      # (the variable: A/B runs of bench.py)

    def _prepared(self, gt: torch.Tensor, fmt: int):
        """(tensor, format) the tile kernel should read for this frame: the prepared copy, made when the
        frame (address, version counter, format) differs from the one prepared last"""
        if not self.prepare_targets:
            return gt, fmt
        key = (gt.data_ptr(), gt._version, fmt)
        if getattr(self, "_tiled_key", None) != key:
            lib = self._cdll
            nbytes = int(lib.gs2d_fit_target_bytes(self.H, self.W))
            buf = getattr(self, "_tiled", None)
            if buf is None or buf.numel() * 4 != nbytes:
                buf = self._tiled = torch.empty((nbytes // 4,), dtype=torch.float32, device=self.device)
            with self._C._guard(self.device):
                self._lib.check(lib.gs2d_fit_target_prepare(self.H, self.W, gt.data_ptr(), fmt, buf.data_ptr(),
                                                            torch.cuda.current_stream(self.device).cuda_stream),
                                "fit_target_prepare")
            self._tiled_key = key
            self._tiled_src = gt          # keeps the address from being recycled while the key is live
        return self._tiled, self._lib.TARGET_TILED

    def train_iters(self, gt_image: torch.Tensor, iters: int = 1, **dbg):
        gt, fmt = self._target(gt_image)
        if gt.device != self.device:
            raise RuntimeError("target frame must live on the model's device")
        gt, fmt = self._prepared(gt, fmt)
        with self._C._guard(self.device):
            for _ in range(int(iters)):
                self._issue(gt, fmt, **dbg)

    def train_until(self, gt_image: torch.Tensor, target_step: int, check_every: int = 0, max_recoveries: int = 8):
        """Iterate until `target_step` iterations have REALLY happened, recovering from workspace
        overflows on the way (GS2DOverflow: the model rewinds itself, regrows / switches its
        workspace, and the lost iterations are issued again).  check_every: look at the device every
        so many iterations (0: only at the end -- one host wait per call, like train_iters + loss).
        Returns the loss of the last iteration."""
        recoveries = 0
        while True:
            try:
                while self.step_index < target_step:
                    k = target_step - self.step_index
                    if check_every > 0:
                        k = min(k, check_every)
                    self.train_iters(gt_image, k)
                    if check_every > 0 and self.step_index < target_step:
                        self.loss()
                return self.loss()
            except GS2DOverflow:
                recoveries += 1
                if recoveries > max_recoveries:
                    raise

    def _read_report(self, step: int):
        """(loss_sum, num_intersects, overflow_step, max_per_tile) a step wrote into the pinned ring;
        the caller has already waited for that step"""
        slot = step % self._lib.FIT_LOSS_SLOTS
        rec = self._report_np[slot]
        if int(rec[5]) != step:
            raise ValueError(f"step {step}: its report slot has been recycled (or the step has not finished)")
        return float(self._report_f64[slot, 0]), int(rec[2]), int(rec[3]), int(rec[4])

    def status(self):
        """(num_intersects, overflow_step, max_per_tile) of the last issued step; synchronises"""
        torch.cuda.current_stream(self.device).synchronize()
        if self.step_index == 0:
            return 0, 0, 0
        _, i, o, m = self._read_report(self.step_index)
        return i, o, m

    def loss(self, step: int | None = None, wait: bool = True) -> float:
        """mean squared error of `step` (default: the last one).  The step's last kernel wrote it
        into pinned host memory, so this only waits for the stream (wait=False: the caller already
        did, e.g. on an event).  A step's slot is recycled FIT_LOSS_SLOTS steps later."""
        step = self.step_index if step is None else int(step)
        if wait:
            torch.cuda.current_stream(self.device).synchronize()
        loss_sum, _i, overflow, _m = self._read_report(step)
        if overflow != 0:
            raise GS2DOverflow(self._recover(overflow))
        return loss_sum / (3.0 * self.H * self.W)

    def _recover(self, overflow_step: int) -> str:
        """steps >= |overflow_step| were no-ops (parameters and optimizer state untouched); the
        optimizer clock goes back with them (nothing but plain steps since the last control sync).
        Positive: the intersections outgrew `cap` -> double it.  Negative: slot mode met a tile with
        more than 256 Gaussians -> tile ranges from here on (which 256 the reference renders of such
        a tile needs all of its ids)."""
        # the no-op steps still in flight write their reports too: let them finish, then forget every report
        # (the steps are about to be issued again under the same numbers; fit_stream polls the ring)
        torch.cuda.current_stream(self.device).synchronize()
        self._report_np[:, 5] = -1
        self.step_index = abs(overflow_step) - 1
        it0, t0, fresh0, frozen0 = self._ckpt
        d = max(self.step_index - it0, 0)
        self.clock.step, self.clock.fresh, self.clock.lr_frozen = t0 + d, fresh0 and d == 0, frozen0
        if overflow_step < 0:
            self.slots = False
            self.cap = self._cap_ranges
            what = "a tile outgrew its 256 slots: the model switched to tile ranges"
        else:
            self.cap *= 2
            self._cap_ranges = max(self._cap_ranges, self.cap)
            what = f"intersections exceeded the workspace capacity: the workspace was regrown to {self.cap}"
        self._alloc()
        return (f"step {abs(overflow_step)}: {what} and the model rewound to step {self.step_index}: "
                f"re-issue the remaining iterations")

    def train_iter(self, gt_image: torch.Tensor):
        self.train_iters(gt_image, 1)
        loss = self.loss()
        return loss, 10 * math.log10(1.0 / loss)

    def _await_report(self, step: int, stream: int):
        """the step's report lands in pinned host memory with the step's last kernel: poll it
        (cheaper and sooner than an event wait); falls back to a stream sync so that a failed
        launch surfaces as an error instead of a hang"""
        rec = self._report_np[step % self._lib.FIT_LOSS_SLOTS]
        spins = 0
        while rec[5] != step:
            spins += 1
            if spins > 2_000_000:
                torch.cuda.current_stream(self.device).synchronize()
                if rec[5] != step:
                    raise RuntimeError(f"step {step} finished without reporting")

    def _frame_pipe(self, frame_bytes: int, depth: int, fmt: int | None = None):
        """the library's frame pipe (include/gs2d.h) and its ring of device slots, created once per
        model and frame size (a copy stream + 2*depth events: ~0.1 ms that a short stream of frames
        would otherwise pay on every call)"""
        prep = fmt if (fmt is not None and self.prepare_targets) else None
        key = (int(frame_bytes), int(depth), prep)
        if self.__dict__.get("_pipe_key") != key:
            self._drop_frame_pipe()
            ct = self._ctypes
            slots = torch.empty((depth, frame_bytes), dtype=torch.uint8, device=self.device)
            pipe = ct.c_void_p()
            with self._C._guard(self.device):
                self._lib.check(self._cdll.gs2d_frame_pipe_create(slots.data_ptr(), frame_bytes, depth, ct.byref(pipe)),
                                "frame_pipe_create")
            d = self.__dict__
            d["_pipe"], d["_pipe_key"], d["_pipe_slots"] = pipe, key, slots
            if prep is not None:
                # the frames are converted into the tile kernel's own order on the copy stream as well
                nb = int(self._cdll.gs2d_fit_target_bytes(self.H, self.W))
                d["_pipe_prepared"] = torch.empty((depth, nb // 4), dtype=torch.float32, device=self.device)
                self._lib.check(self._cdll.gs2d_frame_pipe_set_prepare(pipe, self.H, self.W, prep,
                                                                        d["_pipe_prepared"].data_ptr()), "frame_pipe_set_prepare")
        return self.__dict__["_pipe"]

    def _drop_frame_pipe(self):
        d = self.__dict__                  # (plain attributes; also safe while the interpreter shuts down)
        pipe = d.get("_pipe")
        if pipe is not None:
            try:
                torch.cuda.current_stream(self.device).synchronize()
                self._cdll.gs2d_frame_pipe_destroy(pipe)
            except Exception:   # noqa: BLE001 -- interpreter shutdown
                pass
        d["_pipe"], d["_pipe_key"], d["_pipe_slots"], d["_pipe_prepared"] = None, None, None, None

    def __del__(self):
        try:
            self._drop_frame_pipe()
        except Exception:   # noqa: BLE001
            pass

    def fit_stream(self, host_frames, prefetch: int = 6, lag: int = 2):
        """One iteration per frame of `host_frames` (an iterable of PINNED host targets of one
        format: raw I420 [H*3/2,W], uint8 RGB [H,W,3] or fp32 [1,3,H,W]); yields (loss, psnr) of
        every step, in order.  Frames travel through the library's frame pipe (include/gs2d.h):
        the copy of frame i+1 runs on the pipe's stream under step i, and a step's loss arrives in
        pinned memory with the step itself.  lag: how many steps may be in flight when a result is
        handed back -- lag=2 (default) keeps two steps queued behind the one whose result is awaited, so the
        host's work hides under the device's; lag=0 is strictly synchronous.  prefetch: frames copied ahead
        (device slots of the pipe).  A 1080p I420 frame is 3.1 MB = 60 us of PCIe time, more than a fit step
        takes: the stream is bound by the host link, and the defaults (6 / 2) are what keeps the link busy
        (measured at C2: 2/1 79.9, 3/1 71.1, 4/2 67.3, 6/2 62.9 us per step against 59.9 us for the copy alone;
        profiles/r2_e2e_sweep.txt).
        A workspace overflow (GS2DOverflow: a tile beyond its 256 slots, more intersections than the
        capacity) is absorbed: the model rewinds, and the frames of the steps that became no-ops --
        still referenced here -- are issued again on the regrown / switched workspace, in order."""
        import itertools
        dev = self.device
        ct, lib, check = self._ctypes, self._cdll, self._lib.check
        it = iter(host_frames)
        first = next(it, None)
        if first is None:
            return
        if not first.is_pinned():
            raise ValueError("fit_stream wants pinned host frames (tensor.pin_memory())")
        fmt = self._target(first)[1]
        frame_bytes = first.numel() * first.element_size()
        depth = max(2, min(int(prefetch), 8))
        pipe = self._frame_pipe(frame_bytes, depth, fmt)
        if self.prepare_targets:
            fmt_in, fmt = fmt, self._lib.TARGET_TILED       # acquire() hands out the converted copy
        else:
            fmt_in = fmt
        slot_out, ptr_out = ct.c_int(), ct.c_void_p()
        stream = torch.cuda.current_stream(dev).cuda_stream
        queued, inflight = [], []        # (slot, host frame) copied but not consumed; (step, host frame) issued
        it = itertools.chain((first,), it)

        def push(frame):
            if frame.numel() * frame.element_size() != frame_bytes or not frame.is_pinned():
                raise ValueError("fit_stream: every frame must be pinned and of the first frame's size")
            rc = lib.gs2d_frame_pipe_push(pipe, frame.data_ptr(), ct.byref(slot_out))
            if rc:
                check(rc, "frame_pipe_push")
            queued.append((slot_out.value, frame))

        def result(step):
            self._await_report(step, stream)
            loss = self.loss(step, wait=False)
            return loss, 10 * math.log10(1.0 / max(loss, 1e-20))

        def refill():
            with self._C._guard(dev):
                while len(queued) < depth:
                    nxt = next(it, None)
                    if nxt is None:
                        break
                    push(nxt)

        recoveries = 0
        refill()
        while queued or inflight:
            try:
                if queued:
                    slot, frame = queued.pop(0)
                    with self._C._guard(dev):
                        rc = lib.gs2d_frame_pipe_acquire(pipe, slot, stream, ct.byref(ptr_out))
                        if rc:
                            check(rc, "frame_pipe_acquire")
                        self._issue(ptr_out.value, fmt, stream=stream)
                        rc = lib.gs2d_frame_pipe_release(pipe, slot, stream)
                        if rc:
                            check(rc, "frame_pipe_release")
                        inflight.append((self.step_index, frame))   # the frame stays alive until its step reported
                        nxt = next(it, None)
                        if nxt is not None:                          # its copy overlaps the step just issued
                            push(nxt)
                while len(inflight) > (lag if queued else 0):
                    res = result(inflight[0][0])
                    inflight.pop(0)
                    yield res
            except GS2DOverflow:
                # loss() has rewound the model and re-made its workspace; the steps still in `inflight` were
                # no-ops (the first of them is the one that overflowed): their frames, then the frames already
                # copied but not consumed, go round again through a fresh pipe
                recoveries += 1
                if recoveries > 8:
                    raise
                redo = [fr for _st, fr in inflight] + [fr for _sl, fr in queued]
                inflight.clear()
                queued.clear()
                self._drop_frame_pipe()
                pipe = self._frame_pipe(frame_bytes, depth, fmt_in)
                it = itertools.chain(redo, it)
                refill()

    @property
    def last_num_intersects(self) -> int:
        return self.status()[0]

    @property
    def get_features(self):
        """colours as rendered: _features_dc * rgb_W (GaussianSplats_Represent.py:62-63)"""
        return self._features_dc if self.rgb_W is None else (self._features_dc * self.rgb_W).contiguous()

    def frame_state(self) -> dict:
        """what the reference keeps of a fitted frame and hands to the next one
        (train_video_Represent.py:107-114): _xyz, _cholesky and the weighted colours"""
        return {"_xyz": self._xyz.detach().clone(), "_cholesky": self._cholesky.detach().clone(),
                "_features_dc": self.get_features.detach().clone()}

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
        _p8, p_xyz, p_chol, p_feat = self._param_ptrs()
        colors = None if self.rgb_W is None else self.get_features      # weighted colours: a tensor of their own
        with self._C._guard(self.device):
            self._lib.check(self._lib.load().gs2d_render_chol(
                self.n, self.cap, self.H, self.W, p_xyz, p_chol,
                p_feat if colors is None else colors.data_ptr(), None, 1 | 8 | (4 if self.all_per_tile else 0),
                self._render_work.data_ptr(),
                self._render_work.numel(),
                out.data_ptr(), 1, 1, None, None, torch.cuda.current_stream(self.device).cuda_stream), "render_chol")
        return out


def _intersections_per_tile_row(xyz_raw: torch.Tensor, chol_raw: torch.Tensor, H: int, W: int):
    """[tile rows] intersections of every tile row for the model's current parameters, as integer torch
    ops on the model's device (identical on every rank): the projection gives radii, the tile bbox of
    helpers.cuh:11-43 of the reference gives the rows and the width each Gaussian adds to them"""
    from .project_gaussians_2d import project_gaussians_2d
    tx, ty = (W + 15) // 16, (H + 15) // 16
    bound = torch.tensor([0.5, 0.0, 0.5], device=xyz_raw.device)
    with torch.no_grad():
        xys, _d, radii, _c, _n = project_gaussians_2d(torch.tanh(xyz_raw).contiguous(), (chol_raw + bound).contiguous(),
                                                     H, W, (tx, ty, 1))
        r = radii.float() / 16
        cx, cy = xys[:, 0] / 16, xys[:, 1] / 16
        x0 = (cx - r).trunc().clamp(0, tx); x1 = (cx + r + 1).trunc().clamp(0, tx)
        y0 = (cy - r).trunc().clamp(0, ty).long(); y1 = (cy + r + 1).trunc().clamp(0, ty).long()
        width = ((x1 - x0).clamp(min=0) * (radii > 0)).double()
        diff = torch.zeros(ty + 1, dtype=torch.float64, device=xyz_raw.device)
        diff.index_add_(0, y0, width)
        diff.index_add_(0, y1, -width)
        return [float(v) for v in diff.cumsum(0)[:ty].round().cpu()]


def split_tile_rows(tile_rows: int, world: int, weights=None):
    """[(row_begin, row_end)] per rank: contiguous bands of tile rows.  weights (one number per
    tile row, e.g. intersections per row) balances the bands by work instead of by rows: the
    partition minimises the heaviest band (binary search on the bound + greedy packing)."""
    if world < 1 or tile_rows < world:
        raise ValueError("need at least one tile row per rank")
    if weights is None:
        edges = [round(r * tile_rows / world) for r in range(world + 1)]
        return [(edges[r], edges[r + 1]) for r in range(world)]
    w = [max(float(x), 0.0) for x in weights]
    if len(w) != tile_rows:
        raise ValueError("one weight per tile row")

    def pack(bound):
        bands, start, acc = [], 0, 0.0
        for row, x in enumerate(w):
            if row > start and acc + x > bound:
                bands.append((start, row))
                start, acc = row, 0.0
            acc += x
        bands.append((start, tile_rows))
        return bands

    lo, hi = max(w), sum(w)
    for _ in range(60):
        mid = 0.5 * (lo + hi)
        if len(pack(mid)) <= world:
            hi = mid
        else:
            lo = mid
    bands = pack(hi)
    while len(bands) < world:   # fewer bands than ranks: split the band with the most rows
        i = max(range(len(bands)), key=lambda k: bands[k][1] - bands[k][0])
        a, b = bands[i]
        m = (a + b) // 2
        bands[i:i + 1] = [(a, m), (m, b)]
    return bands


class SpatialFrameFit(FusedFrameFit):
    """One frame split across the ranks of a process group by bands of tile rows (SURVEY 8e,
    BASELINE config 5: 4K+ frames).  Every rank holds all Gaussians and the whole target frame;
    it bins and rasterises only its band, one all-reduce sums the packed gradient rows (and, in
    the same buffer, the partial losses), then every rank applies the identical update.  The
    reference has nothing like it (single `cuda:0`).

    bands: list of (row_begin, row_end) per rank (default: equal split); `emulate_world` runs all
    bands in this one process on one GPU and sums in place of the all-reduce (tests)."""

    path_name = "tile-row bands per rank; scatter-pack + tile kernel on the band, gradient rows summed across ranks, replicated update"

    def __init__(self, H, W, num_points, device, group=None, bands=None, emulate_world: int = 0,
                 exchange: str = "nccl", slots: bool = False, own_barrier: bool = True, balance: bool = False,
                 remote_scatter: bool = False, push: Optional[bool] = None, **kw):
        """exchange: "nccl" -- one all-reduce call on the gradient rows between the two phases;
        "peer" -- the rows live in symmetric (peer-mapped) memory and the update kernel itself sums
        them over NVLink in rank order after a device-side barrier: no collective kernel, no extra
        pass over the rows, bit-identical results on every rank by construction."""
        sharded = exchange == "sharded"
        if slots and not sharded:
            raise NotImplementedError("tile slots are wired into the tile-row split for exchange='sharded' only")
        super().__init__(H, W, num_points, device, slots=sharded, **kw)
        import torch.distributed as dist
        self._dist = dist
        if self.rgb_W is not None:
            raise NotImplementedError("the prune / add control is not wired into the tile-row split yet")
        self.group = group
        self.own_barrier = bool(own_barrier)     # False: torch's symmetric-memory barrier between two host calls
        # sharded update only: the update kernel writes the next step's records straight into the slots of
        # the rank that rasterises each tile (gs2d_grad_exchange.work_peers) instead of every rank
        # re-projecting all Gaussians to find its band's
        self.remote_scatter = bool(remote_scatter)
        # sharded update only: the tile kernels add every pair's gradient straight into the OWNER rank's buffer
        # (gs2d_grad_exchange.push_per) instead of the update kernel reading every rank's rows
        # Default (None): on.  Measured on C5 (profiles/r2_mg{2,4,8}_bench.json, r2_ab_red_v4.txt): with 16-byte
        # vector reductions over NVLink pushing wins at every rank count (2: 0.129 vs 0.135 ms, 4: 0.095 vs
        # 0.115, 8: 0.084 vs 0.110); with scalar ones it lost at two ranks.
        self.push = bool(push) if push is not None else None
        self._push_env_off = os.environ.get("GS2D_NO_PUSH") == "1"
        self.emulate_world = int(emulate_world)
        if exchange not in ("nccl", "peer", "nvls", "sharded"):
            raise ValueError("exchange must be 'nccl', 'peer', 'nvls' or 'sharded'")
        if exchange == "nvls" and emulate_world:
            raise ValueError("the switch reduction needs real ranks (no single-GPU emulation)")
        self.exchange = exchange
        if self.emulate_world:
            self.world, self.rank = self.emulate_world, 0
        else:
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.push = (True if self.push is None else self.push) and sharded and not self._push_env_off
        ty = (self.H + 15) // 16
        if bands is None and balance:
            # bands "balanced by intersection count (prefix sum over rows of cnt_t)" -- what a band
            # costs is its intersections plus a fixed cost per tile (DESIGN: about a third of the tile
            # kernel at 15 Gaussians per tile), hence the mean added to every row
            rows = self.intersections_per_tile_row()
            mean = sum(rows) / max(len(rows), 1)
            bands = split_tile_rows(ty, self.world, weights=[r + 0.5 * mean for r in rows])
        self.bands = list(bands) if bands is not None else split_tile_rows(ty, self.world)
        if len(self.bands) != self.world or self.bands[0][0] != 0 or self.bands[-1][1] != ty or any(
                a[1] != b[0] for a, b in zip(self.bands, self.bands[1:])):
            raise ValueError("bands must tile [0, tile_rows) contiguously, one per rank")
        lay = self.layout
        self._grads = self.work[lay.off_grads:lay.off_grads + self.n * 48].view(torch.float32)
        self._hdl = None
        if sharded:
            self._setup_sharded()
        if exchange in ("peer", "nvls", "sharded"):
            self._setup_peer_buffers()

    def intersections_per_tile_row(self):
        return _intersections_per_tile_row(self._xyz.detach(), self._cholesky.detach(), self.H, self.W)

    def _setup_sharded(self):
        """Sharded update: rank r owns Gaussians shards[r] (optimizer state, gradient reduction,
        update) and rasterises bands[r].  Its update kernel stores the new parameters into every
        rank's parameter tensors, and a tile that outgrows its slots is flagged in every rank's
        status words: both move into symmetric (peer-mapped) memory."""
        dev, n, lay, R = self.device, self.n, self.layout, self.world
        # uniform shards: Gaussian i belongs to rank i // per (the tile kernels of a pushed reduction compute the owner)
        per = -(-n // R)
        self._shard_per = per
        self.shards = [(r * per, min((r + 1) * per, n)) for r in range(R)]
        if any(b <= a for a, b in self.shards):
            raise ValueError("need at least one Gaussian per rank")
        lib = self._lib.load()
        if self.emulate_world:
            # one process: the "ranks" share the parameter tensors, each has a workspace of its own
            self._works = [self.work] + [torch.empty_like(self.work) for _ in range(R - 1)]
            with self._C._guard(dev):
                for wk in self._works[1:]:
                    self._lib.check(lib.gs2d_fit_init(n, self.cap, self.H, self.W, wk.data_ptr(), lay.total_bytes, None,
                                                      torch.cuda.current_stream(dev).cuda_stream), "fit_init")
            par = [[self._params8.data_ptr()] * R] * 3        # packed rows: only the first list is read
            status = [wk.data_ptr() + lay.off_status for wk in self._works]
        else:
            import torch.distributed._symmetric_memory as symm
            grp = self.group if self.group is not None else self._dist.group.WORLD
            buf = symm.empty((8 * n,), dtype=torch.float32, device=dev)
            self._par_hdl = symm.rendezvous(buf, grp)
            with torch.no_grad():                              # the packed [n][8] rows move into peer-mapped memory
                self._pack(self._xyz.detach().clone(), self._cholesky.detach().clone(),
                           self._features_dc.detach().clone(), into=buf.view(n, 8))
            wk = symm.empty((lay.total_bytes,), dtype=torch.uint8, device=dev)
            self._work_hdl = symm.rendezvous(wk, grp)
            with self._C._guard(dev):
                self._lib.check(lib.gs2d_fit_init(n, self.cap, self.H, self.W, wk.data_ptr(), lay.total_bytes, None,
                                                  torch.cuda.current_stream(dev).cuda_stream), "fit_init")
            self.work = wk
            self._status = wk[lay.off_status:lay.off_status + 64].view(torch.int32)
            self._loss = wk[lay.off_loss:lay.off_loss + 8 * self._lib.FIT_LOSS_SLOTS].view(torch.float64)
            self._grads = wk[lay.off_grads:lay.off_grads + n * 48].view(torch.float32)
            self._works = [wk]
            bases = [int(p) for p in self._par_hdl.buffer_ptrs]
            par = [bases] * 3                                   # rank q's packed rows
            status = [int(p) + lay.off_status for p in self._work_hdl.buffer_ptrs]
        works = [wk.data_ptr() for wk in self._works] if self.emulate_world else [int(p) for p in self._work_hdl.buffer_ptrs]
        self._route_ptrs = torch.tensor(par + [status, works], dtype=torch.int64, device=dev)   # xyz, chol, feat, status, work
        self._route_ptr = [self._route_ptrs[k].data_ptr() for k in range(5)]
        self._band_edges = torch.tensor([b[0] for b in self.bands] + [self.bands[-1][1]], dtype=torch.int32, device=dev)

    def _setup_peer_buffers(self):
        dev, n = self.device, self.n
        row_bytes = n * 48
        if self.emulate_world:
            # one process, one GPU: every emulated rank gets its own pair of buffers
            self._xbuf = torch.zeros((self.world, 2, n, 12), dtype=torch.float32, device=dev)
            ptrs = [[self._xbuf[r, b].data_ptr() for r in range(self.world)] for b in range(2)]
            self._own = [[self._xbuf[r, b].data_ptr() for b in range(2)] for r in range(self.world)]
        else:
            import torch.distributed._symmetric_memory as symm
            grp = self.group if self.group is not None else self._dist.group.WORLD
            self._xbuf = symm.empty((2, n, 12), dtype=torch.float32, device=dev)
            self._xbuf.zero_()
            self._hdl = symm.rendezvous(self._xbuf, grp)
            base = [int(p) for p in self._hdl.buffer_ptrs]
            ptrs = [[base[r] + b * row_bytes for r in range(self.world)] for b in range(2)]
            self._own = [[base[self.rank] + b * row_bytes for b in range(2)]]
            self._mc = int(self._hdl.multicast_ptr or 0)
            if self.exchange == "nvls" and not self._mc:
                raise RuntimeError("exchange='nvls': this process group has no NVLink-switch multicast (NVLS) mapping")
            # pads of the library's own cross-rank barrier (gs2d_grad_exchange.barrier_pads)
            self._pads = symm.empty((64,), dtype=torch.int32, device=dev)
            self._pads.zero_()
            self._pads_hdl = symm.rendezvous(self._pads, grp)
            self._pad_ptrs = torch.tensor([int(p) for p in self._pads_hdl.buffer_ptrs], dtype=torch.int64, device=dev)
            self._barriers = 0                      # barriers issued so far (same on every rank)
            torch.cuda.current_stream(dev).synchronize()
            self._hdl.barrier(channel=0)            # every rank's buffers are zero before anyone adds
        self._peer_lists = torch.tensor(ptrs, dtype=torch.int64, device=dev)   # [2][world] device pointers
        self._peer_list_ptr = [self._peer_lists[b].data_ptr() for b in range(2)]

    def _exchange_struct(self, rank_slot: int, with_peers: bool):
        """gs2d_grad_exchange for this step: accumulate into buffer step&1, zero the other one"""
        if self.exchange == "nccl":
            return None
        x = self.step_index & 1
        ex = self._lib.GradExchange()
        ex.accumulate = self._own[rank_slot][x]
        ex.peers = self._peer_list_ptr[x] if with_peers else None
        ex.num_peers = self.world if with_peers else 0
        ex.zero = self._own[rank_slot][1 - x]
        ex.multicast = (self._mc + x * self.n * 48) if (with_peers and self.exchange == "nvls") else None
        if self.exchange == "sharded":
            if self.push:      # one buffer: a row is re-zeroed by its owner right after it was read (see gs2d.h)
                x = 0
                ex.accumulate, ex.zero, ex.push_per = self._own[rank_slot][0], None, self._shard_per
            ex.peers, ex.num_peers = self._peer_list_ptr[x], self.world
            ex.shard_begin, ex.shard_end = self.shards[rank_slot if self.emulate_world else self.rank]
            ex.xyz_peers, ex.chol_peers, ex.feat_peers, ex.status_peers = self._route_ptr[:4]
            if self.remote_scatter:
                ex.work_peers, ex.band_edges = self._route_ptr[4], self._band_edges.data_ptr()
        return ex

    def _with_barrier(self, ex):
        """the library issues the step's barriers itself (one host call per step)"""
        count = 2 if self.exchange == "sharded" else 1
        ex.barrier_pads, ex.rank, ex.barrier_epoch = self._pad_ptrs.data_ptr(), self.rank, self._barriers + 1
        self._barriers += count
        return ex

    def _band_call(self, gt, fmt, band, phases, precounted, dbg, ex=None, work=None):
        b1, b2, b3 = self.betas
        opts = self._opts
        opts.adan_step, opts.adan_fresh, opts.skip_update, opts.slots = 0, 0, 0, int(self.slots)   # 0: the step's own clock
        opts.analytic_grad = int(self.analytic_grad)
        work = self.work if work is None else work
        slot = self.step_index % self._lib.FIT_LOSS_SLOTS
        out_img, grads_out, ids_out = dbg.get("out_img"), dbg.get("grads_out"), dbg.get("ids_out")
        _p8, p_xyz, p_chol, p_feat = self._param_ptrs()
        opts.params8 = _p8
        self._lib.check(self._cdll.gs2d_fit_step_band(
            self.n, self.cap, self.H, self.W, p_xyz, p_chol,
            p_feat, self.adan_state.data_ptr(), gt.data_ptr(), fmt, work.data_ptr(),
            self.layout.total_bytes, self.step_index, self.current_lr(), b1, b2, b3, self.eps, self.weight_decay,
            1 if precounted else 0, band[0], band[1], phases,
            None if ex is None else self._ctypes.addressof(ex),
            self._ctypes.addressof(opts),
            None if out_img is None else out_img.data_ptr(),
            None if grads_out is None else grads_out.data_ptr(), None if ids_out is None else ids_out.data_ptr(),
            self._report_ptr + 24 * slot, torch.cuda.current_stream(self.device).cuda_stream), "fit_step_band")

    def _issue_sharded(self, gt, fmt, dbg):
        """GRADS on the band (records of the band from the local parameter copy, tile kernel) |
        barrier | UPDATE of the shard (sums its rows from every rank's buffer, Adan, new parameters
        stored into every rank's tensors) | barrier.  The emulation runs each phase for all "ranks"
        in turn; a phase boundary is then a barrier by construction."""
        P = self._lib
        ranks = range(self.world) if self.emulate_world else (self.rank,)
        precounted = self._precount_valid()             # False: first step, or the parameters were edited
        if self.own_barrier and not self.emulate_world:
            self._band_call(gt, fmt, self.bands[self.rank], self._both_phases, precounted, dbg,
                            self._with_barrier(self._exchange_struct(0, True)))
            self._counted_for = self._versions()
            return
        for ph in (P.FIT_PHASE_GRADS, P.FIT_PHASE_UPDATE):
            for r in ranks:
                slot = r if self.emulate_world else 0
                self._band_call(gt, fmt, self.bands[r], ph, precounted, dbg, self._exchange_struct(slot, True),
                                self._works[slot])
            if not self.emulate_world:
                self._hdl.barrier(channel=0)            # device-side, on the current stream
        self._counted_for = self._versions()

    def gather_optimizer_state(self):
        """sharded update: the optimizer state of a Gaussian is current on its owner only (the
        parameters are current everywhere) -- bring every rank's copy up to date, e.g. before a
        checkpoint or a switch of the exchange"""
        if self.exchange != "sharded" or self.emulate_world:
            return
        for r, (a, b) in enumerate(self.shards):
            src = self._dist.get_global_rank(self.group, r) if self.group is not None else r
            for k in range(4):
                self._dist.broadcast(self.adan_state[k, a:b], src=src, group=self.group)

    def _recover(self, overflow_step: int) -> str:
        if self.exchange != "sharded":
            return super()._recover(overflow_step)
        # a tile outgrew its slots (every rank saw the flag): go on with tile ranges, which only the
        # replicated update has.  EVERY rank must get here (read the loss on all ranks).
        self.gather_optimizer_state()
        self.exchange = "peer"
        msg = super()._recover(overflow_step)
        lay = self.layout
        self._grads = self.work[lay.off_grads:lay.off_grads + self.n * 48].view(torch.float32)
        self._xbuf.zero_()
        if not self.emulate_world:
            torch.cuda.current_stream(self.device).synchronize()
            self._hdl.barrier(channel=0)
        return msg + " (sharded update -> replicated update over peer memory)"

    def _issue(self, gt, fmt, **dbg):
        self.step_index += 1
        G, U = self._lib.FIT_PHASE_GRADS, self._lib.FIT_PHASE_UPDATE
        if self.exchange == "sharded":
            return self._issue_sharded(gt, fmt, dbg)
        if self.emulate_world:
            for r, band in enumerate(self.bands):        # every "rank" adds into its (or the one) gradient buffer
                self._band_call(gt, fmt, band, G, False, dbg, self._exchange_struct(r, False))
            ex = self._exchange_struct(0, True)
            self._band_call(gt, fmt, self.bands[0], U, False, dbg, ex)
            if self.exchange == "peer":                  # the other emulated ranks re-zero their buffers too
                x = self.step_index & 1
                self._xbuf[1:, 1 - x].zero_()
            self._counted_for = None                     # the pre-count covered band 0 only
            return
        band = self.bands[self.rank]
        if self.exchange in ("peer", "nvls") and self.own_barrier:
            self._band_call(gt, fmt, band, G | U, self._precount_valid(), dbg,
                            self._with_barrier(self._exchange_struct(0, True)))
        elif self.exchange in ("peer", "nvls"):
            self._band_call(gt, fmt, band, G, self._precount_valid(), dbg, self._exchange_struct(0, False))
            self._hdl.barrier(channel=0)                 # device-side, on the current stream: all bands are in
            self._band_call(gt, fmt, band, U, False, dbg, self._exchange_struct(0, True))
        else:
            self._band_call(gt, fmt, band, G, self._precount_valid(), dbg)
            self._dist.all_reduce(self._grads, op=self._dist.ReduceOp.SUM, group=self.group)
            self._band_call(gt, fmt, band, U, False, dbg)
        self._counted_for = self._versions()

    @torch.no_grad()
    def forward(self):
        raise NotImplementedError("render a banded model through FusedFrameFit.forward on one rank")
