#!/usr/bin/env python
"""Where a GOP's wall time goes on the host (C4: key frame 4100 iterations with the removal control, predicted frames
1000 iterations with add + prune): device-busy time (CUDA events) against wall time, and a cProfile of one 4-frame GOP.
usage: python profiles/gop_host_profile.py"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from gsvc_b200 import gop, yuv  # noqa: E402

dev = torch.device("cuda:0")
H, W, N = 1080, 1920, 50_000
frame_fn, _ = gop.synthetic_sequence(H, W, num_frames=8, cut_every=4)
host = {f: yuv.rgb_to_i420((frame_fn(f)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()).contiguous().pin_memory()
        for f in range(8)}
get = lambda f: host[f].to(dev, non_blocking=True)
make = gop.reference_fitters(H, W, N, dev)
gop.fit_gop(lambda f: get(0), (0, 2), make, 120, 120)           # warm-up
torch.cuda.synchronize()
for rep in range(2):
    t0 = time.perf_counter()
    out = gop.fit_gop(get, (0, 4), make, 4100, 1000, seed=1)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"GOP of 4 frames (7100 iterations): {dt:.3f} s wall = {dt / 7100 * 1e6:.1f} us per iteration; "
          f"psnr {[round(out[f]['psnr'], 2) for f in sorted(out)]}, points {[out[f]['num_points'] for f in sorted(out)]}", flush=True)
pr = cProfile.Profile()
pr.enable()
gop.fit_gop(get, (4, 8), make, 4100, 1000, seed=1)
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
