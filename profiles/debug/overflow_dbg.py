import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from gsvc_b200 import synth
from gsvc_b200.fit import FusedFrameFit, GS2DOverflow
cuda = torch.device("cuda:0")
def setup(slots):
    H, W, n = 64, 64, 3000
    init = synth.gaussians(n, seed=2)
    init["_xyz"] = init["_xyz"] * 0.05
    init["_cholesky"] = init["_cholesky"] * 0 + torch.tensor([1.2, 0.0, 1.2])
    gt = synth.target_frame(H, W, seed=2).to(cuda)
    return gt, FusedFrameFit(H, W, n, cuda, init=init, slots=slots, lr=0.05)
gt, b = setup(False)
lb = []
for i in range(8):
    l, _ = b.train_iter(gt); lb.append((round(l, 6), b.status()))
print("b ranges stepwise:", lb)
gt, a = setup(True)
la = []
for i in range(8):
    try:
        l, _ = a.train_iter(gt); la.append((round(l, 6), a.status()))
    except GS2DOverflow as e:
        la.append(("overflow", a.step_index, a.slots))
print("a slots stepwise:", la)
for K in (3, 4, 6):
    gt, c = setup(True)
    try:
        c.train_iters(gt, K); l = c.loss(); print("c K", K, "no overflow", l)
    except GS2DOverflow as e:
        print("c K", K, "overflow ->", c.step_index, c.slots, str(e)[:60])
        out = []
        while c.step_index < 8:
            l, _ = c.train_iter(gt); out.append(round(l, 6))
        print("   continues:", out)
