"""GPU: size-independent properties of the path at BASELINE.json's FULL sizes (C2 1080p / 50k, C5
4K / 200k), where the C oracle is too slow to be the checker: linearity of the accumulated sum in
the colours and of the backward in v_out, conservation of the intersection count through binning,
sortedness / stability of the binned ids, band additivity of the render, and the fused fit step
against the reference pipeline at 4K."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from gsvc_b200 import synth  # noqa: E402
from tests import ref_gpu  # noqa: E402

cuda = torch.device("cuda:0")
FULL = [pytest.param(1080, 1920, 50_000, id="C2-1080p-50k"), pytest.param(2160, 3840, 200_000, id="C5-4K-200k")]


def _x(n, seed=1, trained=False):
    g = synth.gaussians(n, seed=seed, trained_like=trained)
    return {k: g[k].to(cuda) for k in ("means2d", "L", "colors", "opacity")}


def _project(x, H, W):
    import gsvc_b200 as G
    return G.project_gaussians_2d(x["means2d"], x["L"], H, W, ((W + 15) // 16, (H + 15) // 16, 1))


@pytest.mark.parametrize("H,W,n", FULL)
def test_render_is_linear_in_the_colours(H, W, n):
    """out(a*c1 + b*c2) == a*out(c1) + b*out(c2): alpha never depends on the colours (forward.cu:596-609)"""
    import gsvc_b200 as G
    x = _x(n)
    p = _project(x, H, W)
    c1, c2 = x["colors"], torch.rand_like(x["colors"])
    r = lambda c: G.rasterize_gaussians_sum(*p, c, x["opacity"], H, W, 16, 16)
    o1, o2, o12 = r(c1), r(c2), r(0.25 * c1 + 2.0 * c2)
    ref = 0.25 * o1 + 2.0 * o2
    assert (o12 - ref).abs().max().item() <= 1e-5 * max(1.0, ref.abs().max().item())
    # a single unit colour channel renders the per-pixel sum of alphas; it is the same for every channel
    ones = torch.ones_like(c1)
    oa = r(ones)
    assert torch.equal(oa[..., 0], oa[..., 1]) and torch.equal(oa[..., 1], oa[..., 2])


@pytest.mark.parametrize("H,W,n", FULL)
def test_backward_is_linear_in_v_out_and_colour_gradient_is_the_alpha_sum(H, W, n):
    import gsvc_b200 as G
    x = _x(n)
    leaves = {k: x[k].clone().requires_grad_(True) for k in x}
    p = _project(leaves, H, W)
    out = G.rasterize_gaussians_sum(*p, leaves["colors"], leaves["opacity"], H, W, 16, 16)
    g = torch.Generator(device="cuda").manual_seed(3)
    v1 = torch.randn(H, W, 3, device=cuda, generator=g)
    v2 = torch.randn(H, W, 3, device=cuda, generator=g)
    names = ("means2d", "L", "colors", "opacity")

    def grads(v):
        return torch.autograd.grad(out, [leaves[k] for k in names], v, retain_graph=True)

    g1, g2, g12 = grads(v1), grads(v2), grads(0.5 * v1 - 3.0 * v2)
    for nm, a, b, c in zip(names, g1, g2, g12):
        ref = 0.5 * a - 3.0 * b
        err = ((c - ref).abs().max() / ref.abs().max().clamp_min(1e-30)).item()
        assert err <= 2e-4, f"{nm}: {err}"
    # d out / d colour_g = sum over pixels of alpha_g: with v_out = 1 every channel gets the same number,
    # and its total over Gaussians is the total of the image rendered with unit colours
    gc = grads(torch.ones(H, W, 3, device=cuda))[2]
    assert (gc[:, 0] - gc[:, 1]).abs().max().item() <= 1e-4 * gc.abs().max().item()
    with torch.no_grad():
        unit = G.rasterize_gaussians_sum(*[t.detach() for t in p], torch.ones_like(x["colors"]), x["opacity"], H, W)
    tot_img, tot_grad = unit[..., 0].double().sum().item(), gc[:, 0].double().sum().item()
    assert abs(tot_img - tot_grad) <= 1e-4 * tot_img


@pytest.mark.parametrize("H,W,n", FULL)
def test_binning_conserves_intersections_and_is_sorted_and_stable(H, W, n):
    import gsvc_b200 as G
    x = _x(n, trained=(H == 1080))
    xys, depths, radii, conics, nth = _project(x, H, W)
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    tiles = tb[0] * tb[1]
    I, cum = G.compute_cumulative_intersects(nth)
    assert I == int(nth.sum().item()) and int(cum[-1].item()) == I
    isect, gids, isect_sorted, gids_sorted, bins = G.bin_and_sort_gaussians(n, I, xys, depths, radii, cum, tb)
    tile_of = (isect_sorted >> 32)
    assert (tile_of[1:] >= tile_of[:-1]).all(), "keys are not sorted"
    same = tile_of[1:] == tile_of[:-1]
    assert (gids_sorted[1:][same] > gids_sorted[:-1][same]).all(), "ids inside a tile are not ascending (stable sort)"
    # the multiset of (tile, id) pairs is preserved, and every Gaussian appears num_tiles_hit times
    assert torch.equal(torch.bincount(gids_sorted.long(), minlength=n), nth.long())
    assert torch.equal(torch.sort(isect)[0], isect_sorted)
    cnt = (bins[:tiles, 1] - bins[:tiles, 0]).long()
    assert int(cnt.sum().item()) == I
    assert torch.equal(cnt, torch.bincount(tile_of, minlength=tiles)[:tiles])
    lo = bins[:tiles, 0][cnt > 0]
    assert torch.equal(lo.long(), (torch.cumsum(cnt, 0) - cnt)[cnt > 0]), "tile ranges are not the exclusive scan"


def test_fused_fit_step_vs_reference_pipeline_at_4k():
    """BASELINE config 5 through gs2d_fit_step: loss and parameter gradients of the first iteration
    against the reference's kernels + torch glue on the same inputs (1e-4 relative)."""
    if not ref_gpu.available():
        pytest.fail("oracle/_ref/libgsplat_ref.so missing")
    from gsvc_b200.fit import FusedFrameFit
    H, W, n = 2160, 3840, 200_000
    g = synth.gaussians(n, seed=4)
    gt = synth.target_frame(H, W).to(cuda)
    fit = FusedFrameFit(H, W, n, cuda, init=g)
    grads = torch.empty((n, 8), device=cuda)
    fit.train_iters(gt, 1, grads_out=grads)
    loss = fit.loss()
    gt_hwc = gt[0].permute(1, 2, 0).contiguous()

    def vfn(out):
        inside = (out > 0) & (out < 1)
        return (2.0 / out.numel()) * (out.clamp(0, 1) - gt_hwc) * inside

    xyz = g["_xyz"].to(cuda)
    means = torch.tanh(xyz)
    L = (g["_cholesky"] + torch.tensor([0.5, 0.0, 0.5])).to(cuda)
    r = ref_gpu.pipeline(means, L, g["_features_dc"].to(cuda), torch.ones(n, 1, device=cuda), H, W, v_out_fn=vfn)
    ref_loss = torch.nn.functional.mse_loss(r["out_img"].clamp(0, 1), gt_hwc).item()
    assert abs(loss - ref_loss) <= 1e-5 * ref_loss
    assert fit.status()[0] == r["num_intersects"]
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()
    assert rel(grads[:, 0:2], r["v_mean2d"] * (1 - means * means)) <= 2e-4
    assert rel(grads[:, 2:5], r["v_L"]) <= 2e-4
    assert rel(grads[:, 5:8], r["v_colors"]) <= 2e-4
