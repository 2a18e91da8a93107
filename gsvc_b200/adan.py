"""Adan (Xie et al., "Adan: Adaptive Nesterov Momentum Algorithm", 2022) as a torch optimizer.

The reference's fit loop constructs `Adan(params, lr=...)` from its own optimizer.py
(optimizer.py:65-74 of the reference: betas (0.98, 0.92, 0.99), eps 1e-8, weight_decay 0,
max_grad_norm 0, no_prox False, bias-corrected, previous gradient := first gradient at step 1).
This is an independent implementation of the published update for THAT configuration; the
constructor takes the reference's keyword arguments so it can stand in for it, and refuses the
non-default settings it does not implement (gradient clipping, no_prox) instead of ignoring them.

`step()` runs one multi-tensor kernel of libgs2d.so (`gs2d_adan_step`, include/gs2d.h) when every
parameter is a contiguous fp32 CUDA tensor (`fused=None`, the default, decides per step;
`fused=True` insists, `fused=False` never) -- the reference's multi-tensor path is ~20
torch._foreach launches per step (optimizer.py:296-362) and it anticipates an external fused
kernel itself (:365-395).  Otherwise a handful of torch._foreach ops (CPU tensors too: the CPU
tests and trajectories use that path).

Per element, with g the gradient and g' the previous one (g' = g at the first step):
    m <- b1 m + (1-b1) g
    d <- b2 d + (1-b2) (g - g')
    u  = g + b2 (g - g')
    n <- b3 n + (1-b3) u^2
    denom = sqrt(n)/sqrt(1-b3^t) + eps
    p <- (p - lr/(1-b1^t) m/denom - lr b2/(1-b2^t) d/denom) / (1 + lr wd)
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch


class Adan(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0, max_grad_norm=0.0,
                 no_prox=False, foreach: bool = True, fused: Optional[bool] = None, capturable: bool = False):
        if lr < 0 or eps < 0 or not all(0.0 <= b < 1.0 for b in betas):
            raise ValueError("Adan: invalid hyper-parameters")
        if max_grad_norm != 0.0:
            raise NotImplementedError("Adan: max_grad_norm (gradient clipping, optimizer.py:148-163) is not implemented; "
                                      "the reference's fit loop leaves it at 0")
        if no_prox:
            raise NotImplementedError("Adan: no_prox=True (optimizer.py:351-357) is not implemented; the reference's "
                                      "fit loop leaves it False")
        self.fused = fused
        # capturable: the step count also lives in a device int32 (group["step_dev"]) which step() bumps
        # with a torch op and the kernel reads at run time, so that step() can sit inside a captured CUDA
        # graph (needs the fused kernel; lr and the other hyper-parameters are baked in at capture)
        self.capturable = bool(capturable)
        del foreach   # accepted for constructor compatibility: both paths here are multi-tensor
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @staticmethod
    def _fusable(ps, gs) -> bool:
        return all(p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and g.is_cuda
                   and g.dtype == torch.float32 and p.device == ps[0].device for p, g in zip(ps, gs))

    def _step_fused(self, ps, gs, ms, ns, ds, prevs, lr, b1, b2, b3, eps, wd, t, first, step_dev=None):
        from . import _lib
        from . import cuda as _C
        gs = [g if g.is_contiguous() else g.contiguous() for g in gs]
        k = len(ps)
        table = lambda ts: (ctypes.c_void_p * k)(*[x.data_ptr() for x in ts])
        counts = (ctypes.c_int * k)(*[p.numel() for p in ps])
        dev = ps[0].device
        with _C._guard(dev):
            _lib.check(_lib.load().gs2d_adan_step(k, table(ps), table(gs), table(ms), table(ns), table(ds), table(prevs),
                                                  counts, lr, b1, b2, b3, eps, wd, t,
                                                  None if step_dev is None else step_dev.data_ptr(),
                                                  1 if first else 0,
                                                  torch.cuda.current_stream(dev).cuda_stream), "adan_step")

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
            fresh = []
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["m"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["n"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["d"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                new_prev = "prev" not in st or t == 1
                fresh.append(new_prev)
                ps.append(p); gs.append(p.grad); ms.append(st["m"]); ns.append(st["n"]); ds.append(st["d"])
                prevs.append(st.get("prev"))
            if not ps:
                continue
            fused = self.fused
            if fused is None or fused:
                ok = self._fusable(ps, gs) and (all(fresh) or not any(fresh))
                if fused and not ok:
                    raise RuntimeError("Adan(fused=True) needs contiguous fp32 CUDA parameters on one device")
                fused = ok
            if self.capturable and not fused:
                raise RuntimeError("Adan(capturable=True) needs the fused kernel (contiguous fp32 CUDA parameters)")
            if fused:
                first = all(fresh)
                if first:   # the kernel writes prev := g; it only needs somewhere to put it
                    for i, p in enumerate(ps):
                        prevs[i] = self.state[p]["prev"] = torch.empty_like(p, memory_format=torch.contiguous_format)
                step_dev = None
                if self.capturable:
                    step_dev = group.get("step_dev")
                    if step_dev is None:
                        step_dev = group["step_dev"] = torch.full((1,), t - 1, dtype=torch.int32, device=ps[0].device)
                    step_dev.add_(1)      # (inside a captured graph: replayed with every iteration)
                self._step_fused(ps, gs, ms, ns, ds, prevs, lr, b1, b2, b3, eps, wd, t, first, step_dev)
                continue
            for i, p in enumerate(ps):
                if fresh[i]:
                    prevs[i] = self.state[p]["prev"] = p.grad.clone()
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
