"""CPU restatement of the reference's fit iteration around the oracle kernels.

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline and
--impl reference legs).  One step = GaussianVideo_frame.train_iter of the reference
(GaussianSplats_Represent.py:191-207): forward :83-90 (tanh / Cholesky-bound activations,
project_gaussians_2d, rasterize_gaussians_sum, clamp), L2 loss (utils.py:21-26), backward
(oracle kernels + tanh / clamp / mse derivatives as torch autograd defines them), Adan step
(optimizer.py:125-236 defaults: betas (0.98, 0.92, 0.99), eps 1e-8, no weight decay,
previous gradient := first gradient at step 1).  fp32 numpy throughout."""
from __future__ import annotations

import math

import numpy as np

from . import oracle as O

_f = np.float32


class CpuFit:
    def __init__(self, H, W, init, gt_hwc, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, tanh=np.tanh):
        """init: dict with _xyz [n,2], _cholesky [n,3], _features_dc [n,3] (numpy or torch);
        gt_hwc: fp32 [H,W,3] in [0,1]; tanh: the activation (tests pass the device's tanhf so that
        a 1-ulp libm difference, which moves a 1080p centre by ~6e-5 px, does not mask parity)"""
        self.tanh = tanh
        self.H, self.W = int(H), int(W)
        self.p = [np.array(init[k], dtype=_f, copy=True) for k in ("_xyz", "_cholesky", "_features_dc")]
        n = self.p[0].shape[0]
        self.m = [np.zeros_like(a) for a in self.p]
        self.n = [np.zeros_like(a) for a in self.p]
        self.d = [np.zeros_like(a) for a in self.p]
        self.prev = None
        self.t = 0
        self.lr, self.betas, self.eps = lr, betas, eps
        self.opacity = np.ones((n, 1), _f)
        self.gt = np.ascontiguousarray(gt_hwc, dtype=_f)
        self.bound = np.array([[0.5, 0.0, 0.5]], _f)
        self.cores = O.num_threads()
        self.last = None

    def grads(self):
        """forward + backward at the current parameters -> (loss, [g_xyz, g_chol, g_feat], detail)"""
        xyz, chol, feat = self.p
        means = np.ascontiguousarray(self.tanh(xyz), dtype=_f)
        gt = self.gt
        scale = _f(2.0 / gt.size)
        r = O.fit_grads(means, chol + self.bound, feat, self.opacity, self.H, self.W,
                        lambda img: scale * (np.clip(img, 0, 1) - gt) * ((img >= 0) & (img <= 1)))
        loss = float(np.mean((np.clip(r["out_img"], 0, 1).astype(np.float64) - gt) ** 2))
        g = [(r["v_mean2d"] * (1 - means * means)).astype(_f), r["v_L"].astype(_f), r["v_colors"].astype(_f)]
        return loss, g, r

    def step(self):
        loss, grads, r = self.grads()
        self.last = r
        self.t += 1
        t, (b1, b2, b3), lr, eps = self.t, self.betas, self.lr, self.eps
        if self.prev is None:
            self.prev = [g.copy() for g in grads]
        bc1, bc2, bc3s = 1 - b1 ** t, 1 - b2 ** t, math.sqrt(1 - b3 ** t)
        for i, g in enumerate(grads):
            diff = g - self.prev[i]
            self.m[i] = (_f(b1) * self.m[i] + _f(1 - b1) * g).astype(_f)
            self.d[i] = (_f(b2) * self.d[i] + _f(1 - b2) * diff).astype(_f)
            u = g + _f(b2) * diff
            self.n[i] = (_f(b3) * self.n[i] + _f(1 - b3) * u * u).astype(_f)
            denom = np.sqrt(self.n[i]) / _f(bc3s) + _f(eps)
            self.p[i] = (self.p[i] - _f(lr / bc1) * self.m[i] / denom - _f(lr * b2 / bc2) * self.d[i] / denom).astype(_f)
            self.prev[i] = g
        return loss
