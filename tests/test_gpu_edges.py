"""GPU: edge cases of the public API (empty / ragged inputs, capacity growth, non-zero depths,
argument errors, render-only mode, drop-in import) -- the reference's Python-side behaviour in
gsplat/gsplat/rasterize_sum.py:53-70,121-129,182-186."""
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

from gsvc_b200 import synth  # noqa: E402
from tests import ref_gpu  # noqa: E402

cuda = torch.device("cuda:0")


def _x(n, seed=1, trained=False):
    g = synth.gaussians(n, seed=seed, trained_like=trained)
    return {k: g[k].to(cuda) for k in ("means2d", "L", "colors", "opacity")}


def _render(x, H, W, **kw):
    import gsvc_b200 as G
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    p = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
    return p, G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, 16, 16, **kw)


def test_no_intersections_returns_background():
    x = _x(100)
    x["means2d"] = x["means2d"] * 0 + 50.0   # far off-screen: radii > 0, num_tiles_hit == 0
    x["means2d"].requires_grad_(True)
    bg = torch.tensor([0.25, 0.5, 0.75], device=cuda)
    p, out = _render(x, 64, 96, background=bg)
    assert (p[2] > 0).all() and (p[4] == 0).all()
    assert out.shape == (64, 96, 3) and torch.equal(out, torch.ones(64, 96, 3, device=cuda) * bg)
    out.sum().backward() if out.requires_grad else None


def test_single_gaussian_and_tiny_image():
    x = _x(1)
    x["means2d"][:] = 0.0
    p, out = _render(x, 16, 16)
    assert out.shape == (16, 16, 3) and out.max() > 0.5 * x["colors"].max()


def test_return_alpha_is_zero_and_uint8_colors():
    x = _x(500)
    H, W = 64, 64
    p, (out, alpha) = _render(x, H, W, return_alpha=True)
    assert alpha.shape == (H, W) and (alpha == 0).all()
    c8 = (x["colors"] * 255).round().to(torch.uint8)
    import gsvc_b200 as G
    o8 = G.rasterize_gaussians_sum(*p, c8, x["opacity"], H, W)
    of = G.rasterize_gaussians_sum(*p, c8.float() / 255, x["opacity"], H, W)
    assert torch.equal(o8, of)


def test_argument_errors():
    import gsvc_b200 as G
    x = _x(10)
    H = W = 32
    p = G.project_gaussians_2d(x["means2d"], x["L"], H, W, (2, 2, 1))
    with pytest.raises(ValueError):
        G.rasterize_gaussians_sum(p[0].view(-1), p[1], p[2], p[3], p[4], x["colors"], x["opacity"], H, W)
    with pytest.raises(ValueError):
        G.rasterize_gaussians_sum(*p, x["colors"].view(-1), x["opacity"], H, W)
    with pytest.raises(ValueError):
        G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, 8, 8)
    with pytest.raises(NotImplementedError):
        G.rasterize_gaussians_sum(*p, torch.rand(10, 4, device=cuda), x["opacity"], H, W)
    with pytest.raises(AssertionError):
        G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, background=torch.ones(4, device=cuda))
    with pytest.raises(ValueError):
        G.project_gaussians_2d_scale_rot(torch.zeros(0, 2, device=cuda), torch.zeros(0, 2, device=cuda),
                                         torch.zeros(0, 1, device=cuda), H, W, (2, 2, 1))
    with pytest.raises(RuntimeError):
        G.rasterize_gaussians_sum(*[t.cpu() for t in p], x["colors"].cpu(), x["opacity"].cpu(), H, W)


def _reset_capacity(H, W, cap=None):
    from gsvc_b200 import rasterize_sum as RS
    dctx = RS._device_ctx(cuda)
    torch.cuda.synchronize()
    try:
        dctx.poll()
    except RS.CapacityOverflow:
        pass
    dctx.cap.pop((H, W), None)
    dctx.cap_n.pop((H, W), None)
    if cap is not None:
        dctx.cap[(H, W)] = cap
    return dctx


@pytest.mark.parametrize("binning", ["ranked", "onesweep"])
def test_capacity_overflow_regrows_and_matches(binning):
    """options.sync_free = False (and the step-by-step chain): a call whose intersections outgrow the
    buffer is redone with a larger one before it returns"""
    import gsvc_b200
    from gsvc_b200 import rasterize_sum as RS
    x = _x(20_000, trained=True)
    H, W = 256, 256
    with gsvc_b200.options.override(binning=binning, sync_free=False):
        dctx = _reset_capacity(H, W, RS._CAP_QUANTUM)   # far too small on purpose
        p, out_small = _render(x, H, W)
        assert dctx.cap[(H, W)] > RS._CAP_QUANTUM
        p2, out_again = _render(x, H, W)
    assert torch.equal(out_small, out_again)
    if ref_gpu.available():
        r = ref_gpu.pipeline(x["means2d"], x["L"], x["colors"], x["opacity"], H, W)
        assert (out_small - r["out_img"]).abs().max().item() <= 1e-5 * max(1.0, r["out_img"].abs().max().item())


def test_sync_free_overflow_is_loud_then_recovers():
    """default (sync-free): the first call of an image size learns the count; a later call that still
    outgrows its buffer gets a NaN image, the next call raises CapacityOverflow with the buffer
    regrown, and the call after that is right again"""
    from gsvc_b200 import rasterize_sum as RS
    small, big = _x(20_000), _x(20_000, trained=True)      # same count, many times the intersections
    H, W = 256, 256
    _reset_capacity(H, W)
    _, ok = _render(small, H, W)                 # first call: waits once, sizes the buffer for `small` (x4)
    assert torch.isfinite(ok).all()
    RS._device_ctx(cuda).cap[(H, W)] = RS._CAP_QUANTUM   # pretend the hint is far too small for what comes
    _, bad = _render(big, H, W)                  # outgrows it: nothing waits, so nothing can be redone
    torch.cuda.synchronize()
    assert torch.isnan(bad).all()
    with pytest.raises(RS.CapacityOverflow):
        _render(big, H, W)
    _, good = _render(big, H, W)
    torch.cuda.synchronize()
    RS.check_overflow(cuda)
    if ref_gpu.available():
        r = ref_gpu.pipeline(big["means2d"], big["L"], big["colors"], big["opacity"], H, W)
        assert (good - r["out_img"]).abs().max().item() <= 1e-5 * max(1.0, r["out_img"].abs().max().item())


def test_sync_free_forward_never_synchronises_the_host():
    """after the first call of a size, a forward must return while the device is still busy with
    work queued BEFORE it (the reference blocks on .item(), utils.py:117)"""
    x = _x(10_000)
    H, W = 256, 256
    _reset_capacity(H, W)
    _render(x, H, W)
    torch.cuda.synchronize()
    spin = torch.empty(1 << 28, dtype=torch.uint8, device=cuda)
    done = torch.cuda.Event()
    for _ in range(20):
        spin.zero_()                               # ~20 x 40 us of device work ahead of the forward
    _render(x, H, W)
    done.record()
    assert not done.query(), "the forward waited for the device"
    torch.cuda.synchronize()


def test_ranked_and_onesweep_binning_agree_bit_for_bit():
    import gsvc_b200
    H, W = 200, 328
    for trained in (False, True):
        x = _x(6_000, trained=trained)
        for k in x:
            x[k].requires_grad_(k != "opacity")
        res = {}
        for binning in ("ranked", "onesweep"):
            with gsvc_b200.options.override(binning=binning):
                _reset_capacity(H, W)
                _, out = _render(x, H, W)
                v = torch.linspace(-1, 1, out.numel(), device=cuda).view_as(out)
                g = torch.autograd.grad(out, [x["means2d"], x["L"], x["colors"]], v)
                res[binning] = (out.detach(), g)
        assert torch.equal(res["ranked"][0], res["onesweep"][0])
        for a, b in zip(res["ranked"][1], res["onesweep"][1]):
            assert (a - b).abs().max().item() <= 1e-4 * max(b.abs().max().item(), 1e-30)


def test_forward_backward_capture_in_a_cuda_graph():
    """project_gaussians_2d -> rasterize_gaussians_sum -> backward captured once and replayed: same
    bits as the eager call (gradients: same up to the order of the atomics)"""
    import gsvc_b200 as G
    from gsvc_b200 import rasterize_sum as RS
    H, W, n = 256, 256, 10_000
    tb = (16, 16, 1)
    x = _x(n)
    for k in ("means2d", "L", "colors"):
        x[k].requires_grad_(True)
    # separate leaves for the eager reference: a leaf's gradient accumulator remembers the stream it was
    # first used on, and one created on the default stream would drag that stream into the capture
    xe = {k: v.detach().clone().requires_grad_(v.requires_grad) for k, v in x.items()}
    gt = synth.target_frame(H, W)[0].permute(1, 2, 0).contiguous().to(cuda)
    _reset_capacity(H, W)

    def step(x):
        p = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
        out = G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, 16, 16)
        loss = ((out.clamp(0, 1) - gt) ** 2).mean()
        grads = torch.autograd.grad(loss, [x["means2d"], x["L"], x["colors"]])
        return out, loss, grads

    out_e, loss_e, grads_e = step(xe)             # eager (also sizes the buffer)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step(x)                                   # warm-up on a side stream, as torch's graph recipe asks
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out_g, loss_g, grads_g = step(x)
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    RS.check_overflow(cuda)
    assert torch.equal(out_g, out_e.detach())
    assert loss_g.item() == pytest.approx(loss_e.item(), rel=1e-6)
    for a, b in zip(grads_g, grads_e):
        assert (a - b).abs().max().item() <= 1e-4 * max(b.abs().max().item(), 1e-30)
    # new inputs, same graph: the replay renders them
    with torch.no_grad():
        x["colors"].mul_(0.5)
    graph.replay()
    torch.cuda.synchronize()
    with torch.no_grad():
        p = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
        ref = G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], H, W, 16, 16)
    assert torch.equal(out_g, ref)


def test_edited_depths_are_not_taken_for_zero():
    """the zero-depth shortcut is void once somebody writes into the depths tensor"""
    import gsvc_b200 as G
    x = _x(3_000, trained=True)
    H, W = 128, 128
    tb = (8, 8, 1)
    xys, depths, radii, conics, nth = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
    a = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, x["colors"], x["opacity"], H, W)
    depths.add_(torch.rand_like(depths) + 0.1)    # in place: same tensor object, new version
    b = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, x["colors"], x["opacity"], H, W)
    if ref_gpu.available():
        r = ref_gpu.bin_and_sort(3_000, xys, depths, radii, nth, tb)
        ro, _, _ = ref_gpu.rasterize_sum_forward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], xys, conics,
                                                 x["colors"], x["opacity"])
        assert (b - ro).abs().max().item() <= 1e-5 * max(1.0, ro.abs().max().item())
    assert a.shape == b.shape


def test_nonzero_depths_take_the_64bit_key_path():
    """depth decides the order inside a tile (key = tile<<32 | depth bits); GSVC never does this
    but the signature allows it"""
    import gsvc_b200 as G
    if not ref_gpu.available():
        pytest.fail("oracle/_ref missing")
    x = _x(5_000, trained=True)
    for k in x:
        x[k].requires_grad_(k != "opacity")
    H, W = 128, 128
    tb = (8, 8, 1)
    xys, depths, radii, conics, nth = G.project_gaussians_2d(x["means2d"], x["L"], H, W, tb)
    d2 = torch.rand(5_000, device=cuda) + 0.1
    out = G.rasterize_gaussians_sum(xys, d2, radii, conics, nth, x["colors"], x["opacity"], H, W)
    v = torch.randn_like(out)
    out.backward(v)
    with torch.no_grad():
        r = ref_gpu.bin_and_sort(5_000, xys.detach(), d2, radii, nth, tb)
        ro, rTs, rf = ref_gpu.rasterize_sum_forward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], xys.detach(),
                                                    conics.detach(), x["colors"].detach(), x["opacity"])
        rg = ref_gpu.rasterize_sum_backward(H, W, r["gaussian_ids_sorted"], r["tile_bins"], xys.detach(),
                                            conics.detach(), x["colors"].detach(), x["opacity"], rTs, rf, v)
    assert (out.detach() - ro).abs().max().item() <= 1e-5 * max(1.0, ro.abs().max().item())
    rel = ((x["colors"].grad - rg[2]).abs().max() / rg[2].abs().max()).item()
    assert rel <= 1e-4


def test_render_only_mode_matches_training_mode():
    x = _x(10_000)
    H, W = 256, 256
    with torch.no_grad():
        _, a = _render(x, H, W)
    for k in x:
        x[k].requires_grad_(True)
    _, b = _render(x, H, W)
    assert torch.equal(a, b.detach())


def test_repeated_calls_are_deterministic_forward():
    x = _x(10_000)
    outs = [_render(x, 256, 256)[1] for _ in range(3)]
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])


def test_dropin_import_as_gsplat_and_frame_model_step():
    """GSVC's model files import exactly these two names (GaussianSplats_Represent.py:1-2)."""
    import os
    import gsvc_b200
    sys.path.insert(0, os.path.join(os.path.dirname(gsvc_b200.__file__), "dropin"))
    for m in [k for k in sys.modules if k == "gsplat" or k.startswith("gsplat.")]:
        del sys.modules[m]
    from gsplat.project_gaussians_2d import project_gaussians_2d
    from gsplat.rasterize_sum import rasterize_gaussians_sum
    H, W, n = 128, 192, 4000
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    g = synth.gaussians(n)
    _xyz = torch.nn.Parameter(g["_xyz"].to(cuda))
    _chol = torch.nn.Parameter(g["_cholesky"].to(cuda))
    _feat = torch.nn.Parameter(g["_features_dc"].to(cuda))
    bound = torch.tensor([0.5, 0, 0.5], device=cuda).view(1, 3)
    gt = synth.target_frame(H, W).to(cuda)
    opt = torch.optim.Adam([_xyz, _chol, _feat], lr=1e-2)
    losses = []
    for _ in range(30):   # GaussianSplats_Represent.py:83-90,191-207
        opacity = torch.ones(n, 1, device=cuda)
        xys, depths, radii, conics, nth = project_gaussians_2d(torch.tanh(_xyz), _chol + bound, H, W, tb)
        out = rasterize_gaussians_sum(xys, depths, radii, conics, nth, _feat, opacity, H, W, 16, 16,
                                      background=torch.ones(3, device=cuda), return_alpha=False)
        img = torch.clamp(out, 0, 1).view(-1, H, W, 3).permute(0, 3, 1, 2).contiguous()
        loss = torch.nn.functional.mse_loss(img, gt)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        losses.append(loss.item())
    assert losses[-1] < 0.7 * losses[0], losses


def test_graph_replayed_fit_iteration_follows_the_eager_loop():
    """FrameFit.train_iter_graphed (forward + loss + backward + Adan as one graph replay per iteration)
    against FrameFit.train_iter on the same model: same losses, same parameters, step counters in sync"""
    from gsvc_b200.fit import FrameFit
    H, W, n = 128, 192, 4_000
    gt = synth.target_frame(H, W, seed=3).to(cuda)
    a = FrameFit(H, W, n, cuda, seed=3)
    b = FrameFit(H, W, n, cuda, seed=3)
    la, lb = [], []
    for it in range(12):
        l, _ = a.train_iter_graphed(gt)
        la.append(float(l))
        l, _ = b.train_iter(gt)
        lb.append(float(l))
    assert a.iter == b.iter == 12
    assert a.optimizer.param_groups[0]["step"] == b.optimizer.param_groups[0]["step"] == 12
    assert int(a.optimizer.param_groups[0]["step_dev"].item()) == 12
    assert la == pytest.approx(lb, rel=2e-4)
    assert la[-1] < la[0]
    for k in ("_xyz", "_cholesky", "_features_dc"):
        d = (getattr(a, k) - getattr(b, k)).abs()
        assert float(d.median()) <= 1e-6 and float((d > 1e-3).float().mean()) <= 0.01, k
    # a different frame through the same graph
    gt2 = synth.target_frame(H, W, seed=4).to(cuda)
    l2, _ = a.train_iter_graphed(gt2)
    l2b, _ = b.train_iter(gt2)
    assert float(l2) == pytest.approx(float(l2b), rel=2e-4)
