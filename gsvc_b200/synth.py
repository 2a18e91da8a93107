"""Synthetic workloads of BASELINE.json: the reference's own
initialisation of a frame model (GaussianSplats_Represent.py:28-45 of the reference) and a
deterministic smooth target frame.  Pure torch on CPU, seeded; used by bench.py and the tests."""
from __future__ import annotations

import math

import torch

CONFIGS = {
    "C1": dict(H=256, W=256, N=10_000),
    "C2": dict(H=1080, W=1920, N=50_000),
    "C3": dict(H=1080, W=1920, N=50_000),
    "C5": dict(H=2160, W=3840, N=200_000),
}


def gaussians(n: int, seed: int = 1, trained_like: bool = False, dtype=torch.float32):
    """-> dict(_xyz, _cholesky, _features_dc) raw parameters exactly as the reference initialises
    them, plus the activated tensors the two gsplat calls receive: means2d = tanh(_xyz),
    L = _cholesky + [0.5, 0, 0.5], colors = _features_dc (rgb_W = 1), opacity = 1.
    trained_like widens the Cholesky factor (L = 4*rand + bound): radii up to ~25 px, ~10x the
    intersections, some tiles beyond 256 entries (exercises the reference's truncation)."""
    g = torch.Generator().manual_seed(seed)
    _xyz = torch.atanh(2 * (torch.rand(n, 2, generator=g) - 0.5))
    _cholesky = torch.rand(n, 3, generator=g)
    _features_dc = torch.rand(n, 3, generator=g)
    if trained_like:
        _cholesky = _cholesky * 4.0
    bound = torch.tensor([0.5, 0.0, 0.5]).view(1, 3)
    return dict(
        _xyz=_xyz.to(dtype), _cholesky=_cholesky.to(dtype), _features_dc=_features_dc.to(dtype),
        means2d=torch.tanh(_xyz).to(dtype), L=(_cholesky + bound).to(dtype), colors=_features_dc.to(dtype),
        opacity=torch.ones(n, 1, dtype=dtype),
    )


def target_frame(H: int, W: int, seed: int = 1, shift: float = 0.0, pattern: int = 0):
    """gt[c,y,x] = 0.5 + 0.5*sin(2*pi*(f_c*(x+shift)/W + g_c*y/H)) + uniform noise 0.05, clipped to
    [0,1]; returned as [1,3,H,W] (NCHW, what the reference's loss sees) fp32."""
    g = torch.Generator().manual_seed(seed * 7919 + pattern)
    fc = torch.tensor([1.0, 2.0, 3.0]) + pattern
    gc = torch.tensor([2.0, 1.0, 0.5]) + 0.5 * pattern
    y = torch.arange(H, dtype=torch.float32).view(1, H, 1) / H
    x = (torch.arange(W, dtype=torch.float32).view(1, 1, W) + shift) / W
    img = 0.5 + 0.5 * torch.sin(2 * math.pi * (fc.view(3, 1, 1) * x + gc.view(3, 1, 1) * y))
    img = img + 0.05 * (torch.rand(3, H, W, generator=g) - 0.5)
    return img.clamp_(0, 1).unsqueeze(0).contiguous()
