#!/usr/bin/env python
"""Host-time breakdown of one iteration of the unchanged operator loop (FrameFit.train_iter =
GaussianSplats_Represent.py:191-207 on this package's operators) at C2: torch.profiler CPU self
times per op, plus wall / device times of the loop with and without the per-iteration loss read.
usage: python profiles/dropin_host_profile.py [--binning ranked|onesweep] [--sync] [--foreach-adan]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import gsvc_b200  # noqa: E402
from gsvc_b200 import synth  # noqa: E402
from gsvc_b200.fit import FrameFit  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--binning", default="ranked")
ap.add_argument("--sync", action="store_true", help="options.sync_free = False")
ap.add_argument("--foreach-adan", action="store_true")
ap.add_argument("--iters", type=int, default=200)
a = ap.parse_args()
gsvc_b200.options.binning = a.binning
gsvc_b200.options.sync_free = not a.sync
dev = torch.device("cuda:0")
H, W, N = 1080, 1920, 50_000
fit = FrameFit(H, W, N, dev, seed=1)
if a.foreach_adan:
    fit.optimizer.fused = False
gt = synth.target_frame(H, W, seed=1).to(dev)
for _ in range(20):
    fit.train_iter(gt)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.iters):
    fit.train_iter(gt)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / a.iters
print(f"train_iter (loss.item() every iteration): {dt * 1e6:.1f} us/iter = {1 / dt:.0f} iters/s")
with torch.no_grad():
    for _ in range(10):
        fit.forward()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        fit.forward()
    t_host = (time.perf_counter() - t0) / a.iters
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / a.iters
print(f"forward (render): host {t_host * 1e6:.1f} us/frame issued, {dt * 1e6:.1f} us/frame done = {1 / dt:.0f} fps")
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(20):
        fit.train_iter(gt)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=35, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=60))
