import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
sys.argv = [sys.argv[0]]
import bench
from gsvc_b200.fit import FusedFrameFit
dev = torch.device("cuda:0")
H, W, N = 1080, 1920, 50_000
raw = bench.raw_frame(seed=1).pin_memory()
raws = [raw.clone().pin_memory() for _ in range(8)]
K = 300
def run(prefetch, lag, prepare=True, many=False):
    fit = FusedFrameFit(H, W, N, dev, seed=1)
    fit.prepare_targets = prepare
    src = (lambda k: raws[k % 8]) if many else (lambda k: raw)
    for _ in fit.fit_stream((src(k) for k in range(20)), prefetch=prefetch, lag=lag): pass
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in fit.fit_stream((src(k) for k in range(K)), prefetch=prefetch, lag=lag): pass
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / K * 1e6
for prepare in (True, False):
    for prefetch, lag in ((2, 1), (3, 1), (4, 1), (4, 2), (6, 2), (8, 3)):
        print(f"prepare={prepare} prefetch={prefetch} lag={lag}: {run(prefetch, lag, prepare):.1f} us/step   (8 distinct host buffers: {run(prefetch, lag, prepare, True):.1f})")
# the copy alone, back to back on one stream, and with a fit loop running beside it
gt = raw.to(dev)
s2 = torch.cuda.Stream()
dst = torch.empty_like(gt)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(s2):
    a.record()
    for _ in range(100): dst.copy_(raw, non_blocking=True)
    b.record()
torch.cuda.synchronize()
print(f"copy alone: {a.elapsed_time(b) * 10:.1f} us per frame")
fit = FusedFrameFit(H, W, N, dev, seed=1)
fit.train_iters(gt, 20); torch.cuda.synchronize()
with torch.cuda.stream(s2):
    a.record()
    for _ in range(100): dst.copy_(raw, non_blocking=True)
    b.record()
fit.train_iters(gt, 200)
torch.cuda.synchronize()
print(f"copy beside a running fit loop: {a.elapsed_time(b) * 10:.1f} us per frame")
