"""GPU: the options beyond the reference, all off by default.

* N-channel colours and rendering past the first 256 Gaussians of a tile: checked against the
  reference's own N-channel kernels (forward.cu:630-710, backward.cu:864-978), which the reference
  compiles but never binds -- oracle/_ref launches them.  Images 1e-5, gradients 1e-4
  relative (atomic order differs), integers bit-exact.
* Analytic projection gradients: checked against fp64 autograd of the forward chain
  (foward2d.cu:37-47 / :93-101 -> helpers.cuh:45-58), which the reference's own formula fails
  whenever the off-diagonal term is non-zero."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from gsvc_b200 import synth  # noqa: E402
from tests import ref_gpu  # noqa: E402

cuda = torch.device("cuda:0")
IMG_TOL = 1e-5
GRAD_RTOL = 1e-4


def _need_ref():
    if not ref_gpu.available():
        pytest.fail("oracle/_ref/libgsplat_ref.so missing: run oracle/build_ref.sh in the build container")


def _rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _inputs(n, seed=1, trained_like=False, channels=3):
    g = synth.gaussians(n, seed=seed, trained_like=trained_like)
    x = {k: g[k].to(cuda) for k in ("means2d", "L", "opacity")}
    gen = torch.Generator().manual_seed(seed + 100)
    x["colors"] = torch.rand(n, channels, generator=gen).to(cuda) if channels != 3 else g["colors"].to(cuda)
    x["opacity"] = (0.3 + 0.7 * torch.rand(n, 1, generator=gen)).to(cuda)
    return x


def _tile_end_per_pixel(tile_bins, H, W):
    """[H,W] int32: range.y of the pixel's tile -- handed to the reference's N-channel backward as
    final_index, it makes that kernel give a gradient to every entry of the tile (it starts one
    below final_index, backward.cu:921)."""
    tx, ty = (W + 15) // 16, (H + 15) // 16
    hi = tile_bins[: tx * ty, 1].view(ty, tx)
    return hi.repeat_interleave(16, 0).repeat_interleave(16, 1)[:H, :W].contiguous()


def _run_mine(x, H, W, v_out, **opts):
    import gsvc_b200 as G
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    leaves = {k: x[k].clone().requires_grad_(True) for k in ("means2d", "L", "colors", "opacity")}
    with G.options.override(**opts):
        p = G.project_gaussians_2d(leaves["means2d"], leaves["L"], H, W, tb)
        for t in (p[0], p[3]):
            t.retain_grad()
        out = G.rasterize_gaussians_sum(*p, leaves["colors"], leaves["opacity"], H, W, 16, 16)
        (out * v_out).sum().backward()
    return out.detach(), p, leaves


@pytest.mark.parametrize("H,W,n,channels,trained", [
    pytest.param(256, 256, 10_000, 1, False, id="1ch"),
    pytest.param(130, 250, 3_000, 4, True, id="4ch-ragged"),
    pytest.param(256, 256, 40_000, 5, True, id="5ch-over-256-per-tile"),
    pytest.param(1080, 1920, 50_000, 8, False, id="8ch-1080p"),
    pytest.param(128, 128, 2_000, 13, True, id="13ch"),
    pytest.param(128, 128, 2_000, 32, True, id="32ch"),
])
def test_nd_channels_vs_reference_nd_kernels(H, W, n, channels, trained):
    _need_ref()
    x = _inputs(n, seed=3, trained_like=trained, channels=channels)
    v_out = torch.randn(H, W, channels, device=cuda)
    out, p, leaves = _run_mine(x, H, W, v_out, nd_channels=True)
    xys, depths, radii, conics, nth = (t.detach() for t in p)
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    r = ref_gpu.bin_and_sort(n, xys, depths, radii, nth, tb)
    ids, bins = r["gaussian_ids_sorted"], r["tile_bins"]
    if channels == 5:
        assert (bins[:256, 1] - bins[:256, 0]).max().item() > 256, "fixture must exceed 256 per tile"
    ro, rTs, _ = ref_gpu.nd_rasterize_sum_forward(H, W, ids, bins, xys, conics, x["colors"], x["opacity"])
    assert out.shape == (H, W, channels)
    assert (out - ro).abs().max().item() <= IMG_TOL * max(1.0, ro.abs().max().item())
    ref = ref_gpu.nd_rasterize_sum_backward(H, W, ids, bins, xys, conics, x["colors"], x["opacity"], rTs,
                                            _tile_end_per_pixel(bins, H, W), v_out)
    mine = (p[0].grad, p[3].grad, leaves["colors"].grad, leaves["opacity"].grad)
    for nm, a, b in zip(("v_xy", "v_conic", "v_colors", "v_opacity"), mine, ref):
        assert a.shape == b.shape, nm
        assert _rel(a, b) <= GRAD_RTOL, f"{nm}: rel err {_rel(a, b)}"


def test_nd_needs_the_option_and_limits_channels():
    import gsvc_b200 as G
    x = _inputs(100, channels=4)
    p = G.project_gaussians_2d(x["means2d"], x["L"], 64, 64, (4, 4, 1))
    with pytest.raises(NotImplementedError):
        G.rasterize_gaussians_sum(*p, x["colors"], x["opacity"], 64, 64)
    with G.options.override(nd_channels=True):
        with pytest.raises(RuntimeError):
            G.rasterize_gaussians_sum(*p, torch.rand(100, 33, device=cuda), x["opacity"], 64, 64)
    assert G.options.nd_channels is False


def test_untruncated_rgb_vs_reference_nd_kernels_and_default():
    """max_per_tile = 0 renders every Gaussian of a tile (the reference's 3-channel kernel stops at
    256); = 256 through the same entry point is bit-identical to the default kernels."""
    _need_ref()
    import gsvc_b200 as G
    H, W, n = 256, 256, 40_000
    x = _inputs(n, seed=3, trained_like=True)
    v_out = torch.randn(H, W, 3, device=cuda)
    out_def, p_def, l_def = _run_mine(x, H, W, v_out)
    out_all, p_all, l_all = _run_mine(x, H, W, v_out, max_per_tile=0)
    xys, depths, radii, conics, nth = (t.detach() for t in p_all)
    r = ref_gpu.bin_and_sort(n, xys, depths, radii, nth, (16, 16, 1))
    ids, bins = r["gaussian_ids_sorted"], r["tile_bins"]
    cnt = bins[:256, 1] - bins[:256, 0]
    assert cnt.max().item() > 256
    assert (out_all - out_def).abs().max().item() > 1e-3, "the truncation must matter in this fixture"
    ro, rTs, _ = ref_gpu.nd_rasterize_sum_forward(H, W, ids, bins, xys, conics, x["colors"], x["opacity"])
    assert (out_all - ro).abs().max().item() <= IMG_TOL * max(1.0, ro.abs().max().item())
    ref = ref_gpu.nd_rasterize_sum_backward(H, W, ids, bins, xys, conics, x["colors"], x["opacity"], rTs,
                                            _tile_end_per_pixel(bins, H, W), v_out)
    mine = (p_all[0].grad, p_all[3].grad, l_all["colors"].grad, l_all["opacity"].grad)
    for nm, a, b in zip(("v_xy", "v_conic", "v_colors", "v_opacity"), mine, ref):
        assert _rel(a, b) <= GRAD_RTOL, f"{nm}: rel err {_rel(a, b)}"

    # the general kernels with the reference's limit == the default kernels, bit for bit (forward)
    lib, _lib = G._lib.load(), G._lib
    lay = _lib.StateLayout()
    cap = 1 << 20
    _lib.check(lib.gs2d_state_layout_for(n, cap, 16, 16, lay))
    state = torch.empty(lay.total_bytes, dtype=torch.uint8, device=cuda)
    out_nd = torch.empty(H, W, 3, device=cuda)
    fidx_nd = torch.empty(H, W, dtype=torch.int32, device=cuda)
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.gs2d_rasterize_sum_nd_fwd(n, 3, xys.data_ptr(), None, radii.data_ptr(), conics.data_ptr(),
                                             nth.data_ptr(), x["colors"].data_ptr(), x["opacity"].data_ptr(), H, W,
                                             cap, 256, state.data_ptr(), lay.total_bytes, out_nd.data_ptr(),
                                             fidx_nd.data_ptr(), None, None, s), "nd_fwd")
    _, _, rfidx = ref_gpu.rasterize_sum_forward(H, W, ids, bins, xys, conics, x["colors"], x["opacity"])
    assert torch.equal(out_nd, out_def)
    assert torch.equal(fidx_nd, rfidx)
    grads = torch.empty(n, 12, device=cuda)
    v_col = torch.empty(n, 3, device=cuda)
    _lib.check(lib.gs2d_rasterize_sum_nd_bwd(n, 3, H, W, cap, 256, state.data_ptr(), lay.total_bytes,
                                             x["colors"].data_ptr(), fidx_nd.data_ptr(), v_out.data_ptr(),
                                             grads.data_ptr(), v_col.data_ptr(), s), "nd_bwd")
    assert _rel(grads[:, 0:2], p_def[0].grad) <= GRAD_RTOL
    assert _rel(grads[:, 2:5], p_def[3].grad) <= GRAD_RTOL
    assert _rel(v_col, l_def["colors"].grad) <= GRAD_RTOL
    assert _rel(grads[:, 8:9], l_def["opacity"].grad) <= GRAD_RTOL


def _chol_forward64(means, L, H, W):
    a = L[:, 0] ** 2
    b = L[:, 0] * L[:, 1]
    c = L[:, 1] ** 2 + L[:, 2] ** 2
    det = a * c - b * b
    conic = torch.stack([c / det, -b / det, a / det], 1)
    xy = torch.stack([0.5 * W * means[:, 0] + 0.5 * W, 0.5 * H * means[:, 1] + 0.5 * H], 1)
    return xy, conic


def _scalerot_forward64(means, scales, rot, H, W):
    cr, sr = torch.cos(rot[:, 0]), torch.sin(rot[:, 0])
    sx, sy = scales[:, 0], scales[:, 1]
    m00, m01, m10, m11 = cr * sx, -sr * sx, sr * sy, cr * sy    # foward2d.cu:93-97 as project.cu states it
    a = m00 * m00 + m10 * m10
    b = m00 * m01 + m10 * m11
    c = m01 * m01 + m11 * m11
    det = a * c - b * b
    conic = torch.stack([c / det, -b / det, a / det], 1)
    xy = torch.stack([0.5 * W * means[:, 0] + 0.5 * W, 0.5 * H * means[:, 1] + 0.5 * H], 1)
    return xy, conic


def test_analytic_projection_gradients_vs_fp64_autograd():
    """Kernel level.  Convention of the path: the v_conic.y that reaches the projection backward is
    the derivative w.r.t. ONE off-diagonal matrix entry, i.e. half the derivative w.r.t. the conic.y
    parameter (the rasteriser emits 0.5 * v_sigma * dx * dy, backward.cu:829) -- so the loss below
    weights conic.y by 2 * v_conic.y."""
    import gsvc_b200.cuda as C
    n, H, W = 20_000, 540, 960
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    g = torch.Generator().manual_seed(5)
    means = (2 * torch.rand(n, 2, generator=g) - 1)
    L = torch.rand(n, 3, generator=g) * 3 + torch.tensor([0.5, 0.0, 0.5])
    L[:, 1] -= 1.5                                   # both signs of the off-diagonal
    v_xy = torch.randn(n, 2, generator=g)
    v_conic = torch.randn(n, 3, generator=g)
    w64 = v_conic.double() * torch.tensor([1.0, 2.0, 1.0], dtype=torch.float64)
    dm, dL, dvxy, dvc = (t.to(cuda) for t in (means, L, v_xy, v_conic))

    xys, depths, radii, conics, nth = C.project_gaussians_2d_forward(n, dm, dL, H, W, tb, 0.01)
    assert (radii > 0).all()
    m64, l64 = means.double().requires_grad_(True), L.double().requires_grad_(True)
    xy, conic = _chol_forward64(m64, l64, H, W)
    ((xy * v_xy.double()).sum() + (conic * w64).sum()).backward()
    _, gm, gl = C.project_gaussians_2d_backward(n, dm, dL, H, W, radii, conics, dvxy, None, dvc, analytic=True)
    assert _rel(gm.cpu().double(), m64.grad) <= 1e-5
    torch.testing.assert_close(gl.cpu().double(), l64.grad, rtol=2e-3, atol=2e-5 * l64.grad.abs().max().item())
    # ... and the default (reference formula) is NOT that gradient, except for l22
    _, gm_ref, gl_ref = C.project_gaussians_2d_backward(n, dm, dL, H, W, radii, conics, dvxy, None, dvc)
    assert torch.equal(gm_ref, gm) and torch.equal(gl_ref[:, 2], gl[:, 2])
    assert _rel(gl_ref[:, :2].cpu().double(), l64.grad[:, :2]) > 1e-2

    scales = torch.exp(2 * torch.rand(n, 2, generator=g) - 0.5)
    rot = (torch.rand(n, 1, generator=g) - 0.5) * 6
    ds, dr = scales.to(cuda), rot.to(cuda)
    o = C.project_gaussians_2d_scale_rot_forward(n, dm, ds, dr, H, W, tb, 0.01)
    m64, s64, r64 = (t.double().requires_grad_(True) for t in (means, scales, rot))
    xy, conic = _scalerot_forward64(m64, s64, r64, H, W)
    ((xy * v_xy.double()).sum() + (conic * w64).sum()).backward()
    _, gm, gs, gr = C.project_gaussians_2d_scale_rot_backward(n, dm, ds, dr, H, W, o[2], o[3], dvxy, None, dvc,
                                                              analytic=True)
    assert _rel(gm.cpu().double(), m64.grad) <= 1e-5
    torch.testing.assert_close(gs.cpu().double(), s64.grad, rtol=5e-3, atol=5e-5 * s64.grad.abs().max().item())
    torch.testing.assert_close(gr.cpu().double(), r64.grad, rtol=5e-3, atol=5e-5 * r64.grad.abs().max().item())
    _, _, gs_ref, gr_ref = C.project_gaussians_2d_scale_rot_backward(n, dm, ds, dr, H, W, o[2], o[3], dvxy, None, dvc)
    assert _rel(gs_ref.cpu().double(), s64.grad) > 1e-2 or _rel(gr_ref.cpu().double(), r64.grad) > 1e-2


def _dense_render64(means, L, colors, opac, H, W):
    """project (Cholesky) + accumulated-sum rasterisation over ALL pixel x Gaussian pairs in fp64
    torch -- the math of foward2d.cu:37-47, helpers.cuh:45-58, forward.cu:590-609 with autograd."""
    xy, conic = _chol_forward64(means, L, H, W)
    py, px = torch.meshgrid(torch.arange(H, dtype=torch.float64), torch.arange(W, dtype=torch.float64), indexing="ij")
    dx = xy[:, 0].view(-1, 1, 1) - px
    dy = xy[:, 1].view(-1, 1, 1) - py
    cx, cy, cz = (conic[:, k].view(-1, 1, 1) for k in range(3))
    sigma = 0.5 * (cx * dx * dx + cz * dy * dy) + cy * dx * dy
    alpha = torch.clamp(opac.view(-1, 1, 1) * torch.exp(-sigma), max=1.0)
    keep = ((sigma >= 0) & (alpha >= 1.0 / 255.0)).detach()
    return torch.einsum("nhw,nc->hwc", alpha * keep, colors)


def test_analytic_gradient_of_the_whole_chain_vs_dense_fp64_autograd():
    """End to end: with the option on, autograd through project_gaussians_2d ->
    rasterize_gaussians_sum is the true gradient of the rendered image; with the reference's
    formula (default) the Cholesky gradient is not."""
    import gsvc_b200 as G
    H, W, n = 48, 64, 300
    g = torch.Generator().manual_seed(11)
    means = 1.6 * torch.rand(n, 2, generator=g) - 0.8
    L = torch.rand(n, 3, generator=g) * torch.tensor([2.0, 2.0, 2.0]) + torch.tensor([0.8, -1.0, 0.8])
    colors = torch.rand(n, 3, generator=g)
    opac = 0.3 + 0.7 * torch.rand(n, 1, generator=g)
    v_out = torch.randn(H, W, 3, generator=g)
    leaves64 = [t.double().requires_grad_(True) for t in (means, L, colors, opac)]
    (_dense_render64(*leaves64, H, W) * v_out.double()).sum().backward()

    def mine(analytic):
        lv = [t.to(cuda).requires_grad_(True) for t in (means, L, colors, opac)]
        with G.options.override(analytic_projection_grad=analytic):
            p = G.project_gaussians_2d(lv[0], lv[1], H, W, ((W + 15) // 16, (H + 15) // 16, 1))
            out = G.rasterize_gaussians_sum(*p, lv[2], lv[3], H, W, 16, 16)
            (out * v_out.to(cuda)).sum().backward()
        return [t.grad.cpu().double() for t in lv]

    def l2(a, b):
        return ((a - b).norm() / b.norm()).item()

    got = mine(True)
    for nm, a, b in zip(("means2d", "L", "colors", "opacity"), got, (t.grad for t in leaves64)):
        # fp32 + MUFU.EX2 against fp64; a pixel on the alpha >= 1/255 threshold can flip (it moves
        # single entries by up to ~1 % of the largest gradient, not the whole tensor)
        assert l2(a, b) <= 5e-3, f"{nm}: relative L2 err {l2(a, b)}"
        assert _rel(a, b) <= 3e-2, f"{nm}: rel max err {_rel(a, b)}"
    ref_formula = mine(False)
    assert l2(ref_formula[1][:, :2], leaves64[1].grad[:, :2]) > 5e-2
    assert l2(ref_formula[1][:, 2], leaves64[1].grad[:, 2]) <= 5e-3


def test_fused_fit_step_with_analytic_gradient():
    """gs2d_fit_options.analytic_grad: the fused step hands Adan the same parameter gradients as
    autograd through the drop-in operators with the analytic projection backward."""
    import gsvc_b200 as G
    from gsvc_b200.fit import FusedFrameFit
    H, W, n = 256, 256, 10_000
    g = synth.gaussians(n, seed=2)
    gt = synth.target_frame(H, W).to(cuda)
    got = {}
    for analytic in (False, True):
        fit = FusedFrameFit(H, W, n, cuda, init=g, analytic_grad=analytic)
        grads = torch.empty((n, 8), device=cuda)
        fit.train_iters(gt, 1, grads_out=grads)
        got[analytic] = grads.clone()
    xyz = g["_xyz"].to(cuda).requires_grad_(True)
    chol = g["_cholesky"].to(cuda).requires_grad_(True)
    feat = g["_features_dc"].to(cuda).requires_grad_(True)
    with G.options.override(analytic_projection_grad=True):
        p = G.project_gaussians_2d(torch.tanh(xyz), chol + torch.tensor([0.5, 0.0, 0.5], device=cuda), H, W, (16, 16, 1))
        out = G.rasterize_gaussians_sum(*p, feat, torch.ones(n, 1, device=cuda), H, W, 16, 16)
        loss = torch.nn.functional.mse_loss(out.clamp(0, 1).permute(2, 0, 1).unsqueeze(0), gt)
        loss.backward()
    a = got[True]
    assert _rel(a[:, 0:2], xyz.grad) <= 1e-3
    assert _rel(a[:, 2:5], chol.grad) <= 1e-3
    assert _rel(a[:, 5:8], feat.grad) <= 1e-3
    r = got[False]
    # same raster gradients (up to the order of the atomics), different Cholesky formula for l11 / l21 only
    assert _rel(r[:, 0:2], a[:, 0:2]) <= 1e-4 and _rel(r[:, 4], a[:, 4]) <= 1e-4 and _rel(r[:, 5:8], a[:, 5:8]) <= 1e-4
    assert _rel(r[:, 2:4], a[:, 2:4]) > 1e-2


def test_fused_fit_step_renders_every_gaussian_of_a_tile_on_request():
    """FusedFrameFit(max_per_tile=0): tiles with more than 256 Gaussians are rendered and fitted in
    batches of 256.  Checker: the drop-in operators with options.max_per_tile = 0, themselves pinned
    against the reference's N-channel kernels above.  Ids come out in the reference's sorted order
    for the whole tile range (nothing is dropped), the render equals the operators' image, and the
    default (256) differs visibly on this fixture."""
    import gsvc_b200 as G
    from gsvc_b200.fit import FusedFrameFit
    H, W, n = 96, 128, 6_000
    g = synth.gaussians(n, seed=11, trained_like=True)
    gt = synth.target_frame(H, W, seed=11).to(cuda)
    tb = ((W + 15) // 16, (H + 15) // 16, 1)

    xyz = g["_xyz"].to(cuda).requires_grad_(True)
    chol = g["_cholesky"].to(cuda).requires_grad_(True)
    feat = g["_features_dc"].to(cuda).requires_grad_(True)
    with G.options.override(max_per_tile=0):
        p = G.project_gaussians_2d(torch.tanh(xyz), chol + torch.tensor([0.5, 0.0, 0.5], device=cuda), H, W, tb)
        out = G.rasterize_gaussians_sum(*p, feat, torch.ones(n, 1, device=cuda), H, W, 16, 16)
        loss = torch.nn.functional.mse_loss(out.clamp(0, 1).permute(2, 0, 1).unsqueeze(0), gt)
        loss.backward()
    r = ref_gpu.bin_and_sort(n, *(t.detach() for t in (p[0], p[1], p[2], p[4])), tb)
    cnt = r["tile_bins"][: tb[0] * tb[1], 1] - r["tile_bins"][: tb[0] * tb[1], 0]
    assert cnt.max().item() > 256 + 64

    fit = FusedFrameFit(H, W, n, cuda, init=g, slots=False, max_per_tile=0)
    img0 = fit.forward()
    assert (img0[0].permute(1, 2, 0) - out.detach().clamp(0, 1)).abs().max().item() <= 1e-5
    grads = torch.empty((n, 8), device=cuda)
    ids = torch.full((fit.cap,), -7, dtype=torch.int32, device=cuda)
    fit.train_iters(gt, 1, grads_out=grads, ids_out=ids)
    assert fit.status()[1] == 0
    I = r["num_intersects"]
    assert torch.equal(ids[:I], r["gaussian_ids_sorted"])
    assert abs(fit.loss() - loss.item()) <= 1e-5 * loss.item()
    assert _rel(grads[:, 0:2], xyz.grad) <= 2e-4
    assert _rel(grads[:, 2:5], chol.grad) <= 2e-4
    assert _rel(grads[:, 5:8], feat.grad) <= 2e-4

    # a model on tile slots moves to ranges at the first fuller tile and then renders all of it too
    fit_s = FusedFrameFit(H, W, n, cuda, init=g, slots=True, max_per_tile=0)
    fit_s.train_iters(gt, 1)
    with pytest.raises(RuntimeError):
        fit_s.loss()
    assert fit_s.slots is False
    fit_s.train_iters(gt, 1)
    assert abs(fit_s.loss() - loss.item()) <= 1e-5 * loss.item()

    ref256 = FusedFrameFit(H, W, n, cuda, init=g, slots=False)
    ref256.train_iters(gt, 1)
    assert abs(ref256.loss() - loss.item()) > 1e-3 * loss.item()
