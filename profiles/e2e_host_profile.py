#!/usr/bin/env python
"""Where the host's time goes in the end-to-end path (FusedFrameFit.fit_stream at C2, first 200 steps of a fresh
model as bench.py's e2e leg runs them): host time per issued step of the resident loop, per streamed step, and a
cProfile of the streamed loop.  usage: python profiles/e2e_host_profile.py"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

sys.argv = [sys.argv[0]]
import bench  # noqa: E402
from gsvc_b200.fit import FusedFrameFit  # noqa: E402

dev = torch.device("cuda:0")
H, W, N = 1080, 1920, 50_000
raw = bench.raw_frame(seed=1).pin_memory()
gt = raw.to(dev)
K = 200


def fresh(stream):
    fit = FusedFrameFit(H, W, N, dev, seed=1)
    if stream:
        for _ in fit.fit_stream(raw for _ in range(20)):
            pass
    else:
        fit.train_iters(gt, 20)
    torch.cuda.synchronize()
    return fit


for rep in range(3):
    fit = fresh(False)
    t0 = time.perf_counter()
    fit.train_iters(gt, K)
    t_issue = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_done = time.perf_counter() - t0
    print(f"resident train_iters: host {t_issue / K * 1e6:.1f} us per step issued, {t_done / K * 1e6:.1f} us per step done")
    fit = fresh(True)
    t0 = time.perf_counter()
    for _ in fit.fit_stream(raw for _ in range(K)):
        pass
    torch.cuda.synchronize()
    print(f"fit_stream: {(time.perf_counter() - t0) / K * 1e6:.1f} us per step end to end")
fit = fresh(True)
pr = cProfile.Profile()
pr.enable()
for _ in fit.fit_stream(raw for _ in range(K)):
    pass
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr, stream=sys.stdout)
st.sort_stats("tottime").print_stats(16)
