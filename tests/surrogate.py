"""A tiny differentiable stand-in for the two operators, used ONLY to pin the control flow around
them (prune / add schedule, optimizer-state semantics) against the reference's own model class:
tests/golden/make_golden_control.py drives the reference's GaussianVideo_frame with it on CPU,
tests/test_density_cpu.py drives gsvc_b200.fit.FrameFit with it and must reproduce the recorded
trajectory.  It gives every parameter a smooth gradient; it is not the rasteriser."""
import torch


def project(means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh=0.01):
    xys = means2d * torch.tensor([0.5 * img_width, 0.5 * img_height]) + torch.tensor([0.5 * img_width, 0.5 * img_height])
    n = means2d.shape[0]
    return xys, torch.zeros(n), torch.ones(n, dtype=torch.int32), L_elements, torch.ones(n, dtype=torch.int32)


def rasterize(xys, depths, radii, conics, num_tiles_hit, colors, opacity, img_height, img_width, BLOCK_H=16,
              BLOCK_W=16, background=None, return_alpha=False):
    ys, xs = torch.meshgrid(torch.arange(img_height, dtype=torch.float32),
                            torch.arange(img_width, dtype=torch.float32), indexing="ij")
    dx = xs[None] - xys[:, 0, None, None]
    dy = ys[None] - xys[:, 1, None, None]
    a = 0.05 * (1 + conics[:, 0, None, None] ** 2)
    c = 0.05 * (1 + conics[:, 2, None, None] ** 2)
    b = 0.02 * torch.tanh(conics[:, 1, None, None])
    alpha = torch.exp(-(a * dx * dx + c * dy * dy + b * dx * dy))            # [n,H,W]
    return 0.1 * torch.einsum("nhw,nc->hwc", alpha, colors * opacity)   # 0.1: keeps the image inside the clamp
