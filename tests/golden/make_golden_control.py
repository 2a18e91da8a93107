"""Golden trajectories of the reference's prune / add control, from the reference's OWN model class.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_control.py
It imports GaussianSplats_Represent.GaussianVideo_frame and optimizer.Adan from the reference
tree unmodified, on CPU, with the two gsplat operators replaced by tests/surrogate.py (the CUDA
operators cannot run here; the control flow and optimizer semantics being pinned do not depend on
what the renderer computes) and the reference's `utils` module replaced by a stub that provides
the L2 branch of loss_fn (utils.py:21-26; the real module imports cv2 / pytorch_msssim).
Two runs, the two modes train_video_Represent.py uses (:362-366):
  removal:  K-frame, isremoval=True   -> removal_control every interval up to iteration 4000
  density:  P-frame, isdensity=True, initialised from a previous frame's state dict
            -> random add at iteration 1, prune at 500..900, optimizer rebuild at 1000
Recorded per run: the initial state, the number of Gaussians after every iteration, the loss of
every iteration, and parameter snapshots at chosen iterations -> control.npz."""
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import surrogate  # noqa: E402

REF = "/root/reference"


def _install_stubs():
    g = types.ModuleType("gsplat")
    g.__path__ = []
    p = types.ModuleType("gsplat.project_gaussians_2d")
    p.project_gaussians_2d = surrogate.project
    r = types.ModuleType("gsplat.rasterize_sum")
    r.rasterize_gaussians_sum = surrogate.rasterize
    u = types.ModuleType("utils")
    u.F = F

    def loss_fn(pred, target, loss_type="L2", lambda_value=0.7):
        assert loss_type == "L2"
        return F.mse_loss(pred.float(), target.detach().float())

    u.loss_fn = loss_fn
    u.__all__ = ["F", "loss_fn"]
    sys.modules.update({"gsplat": g, "gsplat.project_gaussians_2d": p, "gsplat.rasterize_sum": r, "utils": u})
    sys.path.insert(0, REF)


H = W = 8
SNAP = {"removal": [1, 2, 499, 500, 501, 502, 3999, 4000, 4001, 4002, 4100],
        "density": [1, 2, 3, 499, 500, 501, 900, 901, 1000, 1001, 1002, 1100]}


def target():
    g = torch.Generator().manual_seed(7)
    return torch.rand(1, 3, H, W, generator=g)


def run(mode):
    from GaussianSplats_Represent import GaussianVideo_frame
    torch.manual_seed(3)
    kw = dict(loss_type="L2", opt_type="adan", densification_interval=500 if mode == "removal" else 100,
              iterations=0, H=H, W=W, BLOCK_H=16, BLOCK_W=16, device=torch.device("cpu"), lr=1e-3, quantize=False,
              removal_rate=0.25)
    if mode == "removal":
        m = GaussianVideo_frame(num_points=48, max_num_points=48, isdensity=False, isremoval=True, **kw)
        iters = 4100
    else:
        m = GaussianVideo_frame(num_points=36, max_num_points=48, isdensity=True, isremoval=False, **kw)
        prev = {"_xyz": torch.atanh(2 * (torch.rand(36, 2) - 0.5)), "_cholesky": torch.rand(36, 3),
                "_features_dc": torch.rand(36, 3)}
        sd = m.state_dict()
        sd.update(prev)                     # train_video_Represent.py:65-70
        m.load_state_dict(sd)
        iters = 1100
    out = {f"{mode}.init.{k}": getattr(m, k).detach().numpy().copy() for k in ("_xyz", "_cholesky", "_features_dc", "rgb_W")}
    out[f"{mode}.rng_state"] = torch.get_rng_state().numpy().copy()     # what the add at iteration 1 draws from
    gt = target()
    ns, losses = [], []
    for it in range(1, iters + 1):
        loss, _ = m.train_iter(gt, it)
        ns.append(m._xyz.shape[0])
        losses.append(float(loss.detach()))
        if it in SNAP[mode]:
            for k in ("_xyz", "_cholesky", "_features_dc", "rgb_W"):
                out[f"{mode}.it{it}.{k}"] = getattr(m, k).detach().numpy().copy()
    out[f"{mode}.n"] = np.array(ns, np.int32)
    out[f"{mode}.loss"] = np.array(losses, np.float64)
    print(mode, "n:", ns[0], "->", ns[-1], "loss", losses[0], "->", losses[-1])
    return out


if __name__ == "__main__":
    _install_stubs()
    data = {"gt": target().numpy()}
    for mode in ("removal", "density"):
        data.update(run(mode))
    np.savez_compressed(os.path.join(HERE, "control.npz"), **data)
    print("wrote control.npz:", len(data), "arrays")
