#!/usr/bin/env python
"""The cases compute-sanitizer is run over (one tool per gpurun call; BASELINE config 1 sizes):
 (i)   drop-in forward + backward (both binnings), (ii) the fused step on tile slots and on tile ranges,
 incl. a forced slot overflow and the recovery, (iii) all_per_tile (ranks published into records),
 (iv) the render chain under programmatic dependent launch, from fp32 and from the quantised wire format,
 (v)   the prepared-target conversion, the fused Adan kernel, the prune.
usage: compute-sanitizer --tool memcheck|racecheck|synccheck python profiles/sanitizer_case.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import gsvc_b200 as G  # noqa: E402
from gsvc_b200 import synth  # noqa: E402
from gsvc_b200.fit import FrameFit, FusedFrameFit, GS2DOverflow  # noqa: E402

dev = torch.device("cuda:0")
H, W, n = 256, 256, 10_000
tb = (16, 16, 1)
gt = synth.target_frame(H, W).to(dev)
g = synth.gaussians(n, seed=1)

# (i) operators, both binnings
for binning in ("ranked", "onesweep"):
    with G.options.override(binning=binning):
        x = {k: g[k].to(dev).requires_grad_(k != "opacity") for k in ("means2d", "L", "colors", "opacity")}
        for _ in range(2):
            p = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
            out = G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, 16, 16)
            out.clamp(0, 1).square().mean().backward()
        torch.cuda.synchronize()
print("operators ok")

# (ii) fused step: slots, ranges, u8 / I420 / fp32 targets, prepared and raw
from gsvc_b200 import yuv  # noqa: E402
rgb8 = (gt[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()
frames = [gt, rgb8, yuv.rgb_to_i420(rgb8).contiguous().to(dev)]
for slots in (True, False):
    for prep in (True, False):
        fit = FusedFrameFit(H, W, n, dev, init=g, slots=slots)
        fit.prepare_targets = prep
        for fr in frames:
            fit.train_iters(fr, 2)
        fit.loss()
        with torch.no_grad():
            fit.forward()
torch.cuda.synchronize()
print("fused step ok")

# forced slot overflow + recovery
init = synth.gaussians(3000, seed=2)
init["_xyz"] = init["_xyz"] * 0.05
init["_cholesky"] = init["_cholesky"] * 0 + torch.tensor([1.2, 0.0, 1.2])
fit = FusedFrameFit(64, 64, 3000, dev, init=init, slots=True, lr=0.05)
fit.train_until(synth.target_frame(64, 64, seed=2).to(dev), 4)
print("slot overflow + recovery ok, slots now", fit.slots)

# (iii) every Gaussian of a tile (more than 256), ranks published into the records
fit = FusedFrameFit(64, 64, 3000, dev, init=init, slots=False, max_per_tile=0)
fit.train_iters(synth.target_frame(64, 64, seed=2).to(dev), 2)
fit.loss()
with torch.no_grad():
    fit.forward()
print("all_per_tile ok")

# (iv) render chain (PDL) from fp32 and from the wire format
from gsvc_b200.decode import QuantizedFrame  # noqa: E402
fit = FusedFrameFit(H, W, n, dev, init=g)
books = torch.rand(2, 8, 3, generator=torch.Generator().manual_seed(1)) * torch.tensor([1.0, 0.3]).view(2, 1, 1)
qf = QuantizedFrame.from_state(fit.frame_state(), books, device=dev)
with torch.no_grad():
    for _ in range(2):
        fit.forward()
        qf.render(H, W)
torch.cuda.synchronize()
print("render chain ok")

# (v) operator loop with the fused Adan kernel; prune / add control of the fused path
ff = FrameFit(H, W, n, dev, seed=1)
for _ in range(2):
    ff.train_iter(gt)
fit = FusedFrameFit(H, W, n, dev, seed=1, isremoval=True, densification_interval=2)
fit.train_iters(gt, 5)
fit.loss()
torch.cuda.synchronize()
print("adan + prune ok")
