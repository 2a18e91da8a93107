"""Generate golden vectors from the reference's OWN Python restatement.

Run in the build container only (needs /root/reference, which does not exist on the
GPU box):   python tests/golden/make_golden.py
It imports gsplat/gsplat/_torch_impl.py from the reference tree on CPU (the CUDA
extension loads lazily, so nothing is compiled) and replays the inputs of the
reference's tests:
  gsplat/tests/test_cov2d_bounds.py:9-35       -> cov2d_bounds.npz
  gsplat/tests/test_map_gaussians.py:9-73      -> binning.npz (map_gaussian_to_intersects)
  gsplat/tests/test_get_tile_bin_edges.py:9-81 -> binning.npz (get_tile_bin_edges)
plus one composition of reference pieces for the 2D Cholesky projection
(_torch_impl.compute_cov2d_bounds :197-212 + get_tile_bbox :236-259 on
cov = L L^T, centre = 0.5*W*x + 0.5*W as foward2d.cu:37-47) -> project2d.npz.
The .npz files are committed; tests read only them.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference/gsplat"
sys.path.insert(0, REF)
from gsplat import _torch_impl  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
dev = torch.device("cpu")


def cov2d_bounds_fixture():
    torch.manual_seed(42)
    n = 100
    _covs2d = torch.rand((n, 2, 2), dtype=torch.float32, device=dev)
    covs2d = torch.stack(
        [torch.triu(_covs2d)[:, 0, 0], torch.triu(_covs2d)[:, 0, 1], torch.triu(_covs2d)[:, 1, 1]], dim=-1
    )
    # the reference test feeds the full (non-symmetric) 2x2 to the torch impl, which only
    # reads [0,0], [0,1], [1,1] -- identical to the packed triple.
    conic, radii, mask = _torch_impl.compute_cov2d_bounds(_covs2d)
    np.savez(
        os.path.join(OUT, "cov2d_bounds.npz"),
        covs2d=covs2d.numpy(), conic=conic.numpy(), radii=radii.numpy(), mask=mask.numpy(),
    )
    print("cov2d_bounds:", n, "rows,", int(mask.sum()), "unmasked")


def binning_fixture():
    torch.manual_seed(42)
    num_points = 100
    means3d = torch.randn((num_points, 3), device=dev)
    scales = torch.randn((num_points, 3), device=dev)
    glob_scale = 0.3
    quats = torch.randn((num_points, 4), device=dev)
    quats /= torch.linalg.norm(quats, dim=-1, keepdim=True)
    viewmat = torch.eye(4, device=dev)
    projmat = torch.eye(4, device=dev)
    fx, fy = 3.0, 3.0
    H, W = 512, 512
    clip_thresh = 0.01
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)
    (_cov3d, _xys, _depths, _radii, _conics, _nth, _masks) = _torch_impl.project_gaussians_forward(
        means3d, scales, glob_scale, quats, viewmat, projmat, fx, fy, (H, W), tile_bounds, clip_thresh
    )
    _xys, _depths, _radii, _conics, _nth = (t[_masks] for t in (_xys, _depths, _radii, _conics, _nth))
    n = num_points - int(torch.count_nonzero(~_masks))
    _cum = torch.cumsum(_nth, dim=0, dtype=torch.int32)
    I = int(_cum[-1])
    _depths = _depths.contiguous()
    tmin, tmax = _torch_impl.get_tile_bbox(_xys, _radii, tile_bounds)
    isect, gids = _torch_impl.map_gaussian_to_intersects(n, _xys, _depths, _radii, _cum, tile_bounds)
    isect_sorted, order = torch.sort(isect)
    gids_sorted = torch.gather(gids, 0, order)
    bins = _torch_impl.get_tile_bin_edges(I, isect_sorted)
    np.savez(
        os.path.join(OUT, "binning.npz"),
        H=H, W=W, tile_bounds=np.array(tile_bounds), num_points=n, num_intersects=I,
        xys=_xys.numpy(), depths=_depths.numpy(), radii=_radii.numpy().astype(np.int32),
        num_tiles_hit=_nth.numpy().astype(np.int32), cum_tiles_hit=_cum.numpy(),
        tile_min=tmin.numpy().astype(np.int32), tile_max=tmax.numpy().astype(np.int32),
        isect_ids=isect.numpy(), gaussian_ids=gids.numpy(),
        isect_ids_sorted=isect_sorted.numpy(), gaussian_ids_sorted=gids_sorted.numpy(),
        tile_bins=bins.numpy(),
    )
    print("binning:", n, "gaussians,", I, "intersections; radii>0 all:", bool((_radii > 0).all()))


def project2d_fixture():
    torch.manual_seed(1)
    n, H, W = 2000, 256, 256
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)
    means2d = torch.tanh(torch.atanh(2 * (torch.rand(n, 2) - 0.5)))
    L = torch.rand(n, 3) * torch.tensor([3.0, 2.0, 3.0]) + torch.tensor([0.5, -1.0, 0.5])
    cx = 0.5 * W * means2d[:, 0] + 0.5 * W
    cy = 0.5 * H * means2d[:, 1] + 0.5 * H
    xys = torch.stack([cx, cy], -1)
    cov = torch.zeros(n, 2, 2)
    cov[:, 0, 0] = L[:, 0] * L[:, 0]
    cov[:, 0, 1] = L[:, 0] * L[:, 1]
    cov[:, 1, 0] = cov[:, 0, 1]
    cov[:, 1, 1] = L[:, 1] * L[:, 1] + L[:, 2] * L[:, 2]
    conic, radius, mask = _torch_impl.compute_cov2d_bounds(cov)
    tmin, tmax = _torch_impl.get_tile_bbox(xys, radius, tile_bounds)
    area = (tmax[:, 0] - tmin[:, 0]) * (tmax[:, 1] - tmin[:, 1])
    np.savez(
        os.path.join(OUT, "project2d.npz"),
        H=H, W=W, tile_bounds=np.array(tile_bounds), means2d=means2d.numpy(), L=L.numpy(),
        xys=xys.numpy(), conics=conic.numpy(), radii=radius.to(torch.int32).numpy(),
        num_tiles_hit=area.to(torch.int32).numpy(), mask=mask.numpy(),
    )
    print("project2d:", n, "gaussians, radii", int(radius.min()), "..", int(radius.max()))


if __name__ == "__main__":
    cov2d_bounds_fixture()
    binning_fixture()
    project2d_fixture()
