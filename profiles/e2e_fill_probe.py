#!/usr/bin/env python
"""Fill and drain of FusedFrameFit.fit_stream at the driver's step count (C2, K = 20 after 5 streamed warm-up steps,
as bench.py's e2e leg): device time of the whole loop per step for several (prefetch, lag), host time stamps of the
first and last results, and the same with K = 200.  usage: python profiles/e2e_fill_probe.py"""
import os
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
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for K in (20, 200):
    for prefetch, lag in ((6, 2), (3, 2), (2, 1)):
        best = None
        for rep in range(3):
            fit = FusedFrameFit(H, W, N, dev, seed=1)
            for _ in fit.fit_stream((raw for _ in range(5)), prefetch=prefetch, lag=lag):
                pass
            torch.cuda.synchronize()
            stamps = []
            a.record()
            t0 = time.perf_counter()
            for _ in fit.fit_stream((raw for _ in range(K)), prefetch=prefetch, lag=lag):
                stamps.append(time.perf_counter() - t0)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            rec = (ms / K * 1e3, stamps[0] * 1e6, stamps[1] * 1e6, stamps[-1] * 1e6)
            best = rec if best is None or rec[0] < best[0] else best
        print(f"K {K:3d} prefetch {prefetch} lag {lag}: {best[0]:.1f} us per step (device, whole loop); "
              f"host: first result at {best[1]:.0f} us, second {best[2]:.0f}, last {best[3]:.0f}", flush=True)
