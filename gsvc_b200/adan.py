"""Adan (Xie et al., "Adan: Adaptive Nesterov Momentum Algorithm", 2022) as a torch optimizer.

The reference's fit loop constructs `Adan(params, lr=...)` from its own optimizer.py
(optimizer.py:65-74 of the reference: betas (0.98, 0.92, 0.99), eps 1e-8, weight_decay 0,
no_prox False, bias-corrected, previous gradient := first gradient at step 1).  This is an
independent implementation of the published update with those defaults; `step()` on CUDA
parameters runs one multi-tensor kernel of libgs2d.so (gs2d_adan_step) when `fused=True`,
else a handful of torch._foreach ops (works on CPU tensors too: used by the CPU baseline).

Per element, with g the gradient and g' the previous one (g' = g at the first step):
    m <- b1 m + (1-b1) g
    d <- b2 d + (1-b2) (g - g')
    u  = g + b2 (g - g')
    n <- b3 n + (1-b3) u^2
    denom = sqrt(n)/sqrt(1-b3^t) + eps
    p <- (p - lr/(1-b1^t) m/denom - lr b2/(1-b2^t) d/denom) / (1 + lr wd)
"""
from __future__ import annotations

import math

import torch


class Adan(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0):
        if lr < 0 or eps < 0 or not all(0.0 <= b < 1.0 for b in betas):
            raise ValueError("Adan: invalid hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            b1, b2, b3 = group["betas"]
            t = group["step"] = group.get("step", 0) + 1
            bc1, bc2, bc3s = 1.0 - b1 ** t, 1.0 - b2 ** t, math.sqrt(1.0 - b3 ** t)
            lr, eps, wd = group["lr"], group["eps"], group["weight_decay"]
            ps, gs, ms, ns, ds, prevs = [], [], [], [], [], []
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["m"] = torch.zeros_like(p)
                    st["n"] = torch.zeros_like(p)
                    st["d"] = torch.zeros_like(p)
                if "prev" not in st or t == 1:
                    st["prev"] = p.grad.clone()
                ps.append(p); gs.append(p.grad); ms.append(st["m"]); ns.append(st["n"]); ds.append(st["d"])
                prevs.append(st["prev"])
            if not ps:
                continue
            diff = torch._foreach_sub(gs, prevs)
            torch._foreach_mul_(ms, b1)
            torch._foreach_add_(ms, gs, alpha=1 - b1)
            torch._foreach_mul_(ds, b2)
            torch._foreach_add_(ds, diff, alpha=1 - b2)
            torch._foreach_mul_(diff, b2)
            torch._foreach_add_(diff, gs)            # u
            torch._foreach_mul_(ns, b3)
            torch._foreach_addcmul_(ns, diff, diff, value=1 - b3)
            denom = torch._foreach_sqrt(ns)
            torch._foreach_div_(denom, bc3s)
            torch._foreach_add_(denom, eps)
            torch._foreach_addcdiv_(ps, ms, denom, value=-lr / bc1)
            torch._foreach_addcdiv_(ps, ds, denom, value=-lr * b2 / bc2)
            if wd != 0.0:
                torch._foreach_div_(ps, 1 + lr * wd)
            for prev, g in zip(prevs, gs):
                prev.copy_(g)
        return loss
