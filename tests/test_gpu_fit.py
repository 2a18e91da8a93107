"""GPU parity of the fused fit step / render path (gs2d_fit_step, gs2d_render_chol) against the
CPU oracle (oracle/fit_cpu.py around the C restatement of the reference kernels), through the
C ABI.  Integers bit-exact; images <= 1e-5 (threshold-flip pixels counted separately); gradients
<= 1e-4 of the tensor's max (atomic order differs); parameters after Adan compared directly."""
import ctypes
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda:0")


def _device_tanh(a):
    """CUDA tanhf, so the oracle sees bit-identical activations (np.tanh differs by an ulp here and
    there, which at 1080p shifts a centre by ~6e-5 px and a pixel by more than the 1e-5 image bar)"""
    return torch.tanh(torch.from_numpy(np.ascontiguousarray(a)).to(cuda)).cpu().numpy()


def _oracle_fit(H, W, init, gt_chw):
    from oracle.fit_cpu import CpuFit
    return CpuFit(H, W, {k: init[k].numpy() for k in ("_xyz", "_cholesky", "_features_dc")},
                  gt_chw.permute(1, 2, 0).contiguous().numpy(), tanh=_device_tanh)


def _setup(H, W, n, seed=3, trained_like=False, cap=None, slots=False):
    """slots=False: tile ranges (the layout whose order / ranges compare one to one with the
    reference's sorted arrays); slots=True: the two-kernel iteration on fixed tile slots"""
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit
    init = synth.gaussians(n, seed=seed, trained_like=trained_like)
    gt = synth.target_frame(H, W, seed=seed)      # [1,3,H,W]
    fit = FusedFrameFit(H, W, n, cuda, init=init, cap=cap, slots=slots)
    return init, gt, fit


def _ids_in_reference_order(ids, tile_bins, tiles, slots):
    """the debug id output as one array in the reference's sorted order: ranges are already that;
    slots hold tile t's ids at [t*256, t*256 + count)"""
    ids = ids.cpu().numpy()
    if not slots:
        return ids
    out = np.full_like(ids, -7)
    for t in range(tiles):
        lo, hi = (int(v) for v in tile_bins[t])
        out[lo:hi] = ids[t * 256:t * 256 + (hi - lo)]
    return out


def _rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


GRAD_RTOL = 1e-4          # north_star: gradients within 1e-4 relative of the reference build
ORACLE_GRAD_RTOL = 1e-3   # vs the IEEE CPU restatement: alpha-threshold flips (exp vs ex2.approx) add ~5e-4


def _ref_gpu_param_grads(init, gt, H, W):
    """the reference's own CUDA kernels (oracle/_ref) on the same inputs, chained through the
    activations / loss exactly as autograd would: -> [g_xyz, g_chol, g_feat] numpy, or None"""
    from tests import ref_gpu
    if not ref_gpu.available():
        pytest.fail("oracle/_ref/libgsplat_ref.so missing: run oracle/build_ref.sh in the build container")
    xyz = init["_xyz"].to(cuda)
    means = torch.tanh(xyz)
    L = (init["_cholesky"] + torch.tensor([0.5, 0.0, 0.5])).to(cuda)
    col = init["_features_dc"].to(cuda)
    gt_hwc = gt[0].permute(1, 2, 0).contiguous().to(cuda)

    def vfn(out):
        return (2.0 / out.numel()) * (out.clamp(0, 1) - gt_hwc) * ((out >= 0) & (out <= 1))
    r = ref_gpu.pipeline(means, L, col, torch.ones(xyz.shape[0], 1, device=cuda), H, W, v_out_fn=vfn)
    g = [r["v_mean2d"] * (1 - means * means), r["v_L"], r["v_colors"]]
    return [t.cpu().numpy() for t in g], r


@pytest.mark.parametrize("slots", [False, True])
@pytest.mark.parametrize("H,W,n", [(256, 256, 10_000), (200, 328, 3_000), (1080, 1920, 50_000)])
def test_first_step_grads_binning_loss_and_update(H, W, n, slots):
    init, gt, fit = _setup(H, W, n, slots=slots)
    o = _oracle_fit(H, W, init, gt[0])
    o_loss, o_g, r = o.grads()
    grads = torch.full((n, 8), float("nan"), device=cuda)
    ids = torch.full((fit.cap,), -7, dtype=torch.int32, device=cuda)
    fit.train_iters(gt.to(cuda), 1, grads_out=grads, ids_out=ids)
    assert fit.status()[1] == 0       # no overflow (status now describes the NEXT step's binning)
    # binning of this step: the rendering order of every tile range is bit-exact (the tile ranges
    # themselves are compared in test_render_matches_reference_kernels_and_oracle)
    I = r["num_intersects"]
    ids = torch.from_numpy(_ids_in_reference_order(ids, r["tile_bins"], fit.layout.tiles, slots))
    assert np.array_equal(ids[:I].cpu().numpy(), r["gaussian_ids_sorted"])
    assert (ids[I:] == -7).all()
    assert fit.slots == slots
    # loss and gradients
    loss = fit.loss()
    assert abs(loss - o_loss) <= 2e-6 * max(o_loss, 1e-12) + 1e-9, (loss, o_loss)
    g = grads.cpu().numpy()
    assert np.isfinite(g).all()
    r_g, rr = _ref_gpu_param_grads(init, gt, H, W)
    assert np.array_equal(ids[:I].cpu().numpy(), rr["gaussian_ids_sorted"].cpu().numpy())
    for name, mine, ref, ora in (("xyz", g[:, 0:2], r_g[0], o_g[0]), ("chol", g[:, 2:5], r_g[1], o_g[1]),
                                 ("feat", g[:, 5:8], r_g[2], o_g[2])):
        assert _rel(mine, ref) <= GRAD_RTOL, (name, "vs reference kernels", _rel(mine, ref))
        if H * W <= 256 * 256:
            # the IEEE / no-FMA restatement rounds 0.5*W*x + 0.5*W twice where the device (and the
            # reference build) fuse it: at 1080p centres move by an ulp (6e-5 px), so the direct
            # float comparison with the C oracle is kept to the small configs
            assert _rel(mine, ora) <= ORACLE_GRAD_RTOL, (name, "vs C oracle", _rel(mine, ora))
    # Adan step 1: |update| = lr = 1e-3 for every element with a non-zero gradient (sign-like step),
    # so only elements whose gradient is ~0 (and flips sign with the reduction order) may differ
    o.step()
    for name, mine, ref in (("_xyz", fit._xyz, o.p[0]), ("_cholesky", fit._cholesky, o.p[1]),
                            ("_features_dc", fit._features_dc, o.p[2])):
        d = np.abs(mine.cpu().numpy() - ref)
        assert (d > 2e-5).mean() <= 2e-3 and d.max() <= 2.1e-3, (name, d.max(), (d > 2e-5).mean())


@pytest.mark.parametrize("H,W,n,trained", [(256, 256, 10_000, False), (1080, 1920, 50_000, False),
                                           (1080, 1920, 50_000, True), (200, 328, 3_000, False)])
def test_slot_iteration_follows_the_range_iteration(H, W, n, trained):
    """two kernels per iteration on fixed tile slots (the update kernel scatters the next step's
    records) against three kernels on tile ranges: same losses, same intersection counts, same
    per-tile maxima, step after step; then an edit of the parameters behind the library's back"""
    init, gt, a = _setup(H, W, n, seed=8, trained_like=trained, slots=True)
    _, _, b = _setup(H, W, n, seed=8, trained_like=trained, slots=False, cap=1 << 21)
    gtd = gt.to(cuda)
    for it in range(10):
        la, _ = a.train_iter(gtd)
        lb, _ = b.train_iter(gtd)
        assert la == pytest.approx(lb, rel=2e-5), it
        assert a.status() == b.status(), it                 # (num_intersects, overflow, max_per_tile)
    assert a.slots and a.status()[0] > 0
    for k in ("_xyz", "_cholesky", "_features_dc"):
        d = (getattr(a, k) - getattr(b, k)).abs()
        assert float(d.median()) <= 1e-6 and float((d > 1e-4).float().mean()) <= 0.01, k
    snap = {k: getattr(b, k).detach().clone() for k in ("_xyz", "_cholesky", "_features_dc")}
    snap["_xyz"][::3, 1] += 0.2                             # and moved: the stale scatter must not be used
    with torch.no_grad():
        for m in (a, b):
            for k, v in snap.items():
                getattr(m, k).copy_(v)                      # identical parameters, versions bumped
    a.adan_state.copy_(b.adan_state)
    for it in range(3):
        la, _ = a.train_iter(gtd)
        lb, _ = b.train_iter(gtd)
        assert la == pytest.approx(lb, rel=1e-5), it
        assert a.status() == b.status(), it


def test_slot_overflow_falls_back_to_ranges():
    """a tile with more than 256 Gaussians cannot be rendered from slots (which 256 needs all ids):
    the step is a no-op, the model switches to ranges and continues with identical results"""
    H, W, n = 96, 128, 6_000
    init, gt, a = _setup(H, W, n, seed=11, trained_like=True, slots=True)
    _, _, b = _setup(H, W, n, seed=11, trained_like=True, slots=False)
    gtd = gt.to(cuda)
    before = a._xyz.clone()
    a.train_iters(gtd, 3)
    with pytest.raises(RuntimeError, match="tile ranges"):
        a.loss()
    assert not a.slots and a.step_index == 0 and torch.equal(before, a._xyz)
    a.train_iters(gtd, 3)
    b.train_iters(gtd, 3)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-6)
    assert a.status() == b.status() and a.status()[2] > 256


def test_slot_overflow_in_a_later_step():
    """the overflow is met by the update kernel while it scatters step s+1: step s must be
    complete and valid, step s+1 the first no-op"""
    H, W, n = 64, 64, 3_000
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit
    init = synth.gaussians(n, seed=2)
    init["_xyz"] = init["_xyz"] * 0.05            # everything near the centre: ~190 per tile at the start
    init["_cholesky"] = init["_cholesky"] * 0 + torch.tensor([1.2, 0.0, 1.2])
    gt = synth.target_frame(H, W, seed=2).to(cuda)
    a = FusedFrameFit(H, W, n, cuda, init=init, slots=True, lr=0.05)
    b = FusedFrameFit(H, W, n, cuda, init=init, slots=False, lr=0.05)
    hit = None
    for it in range(1, 200):
        try:
            la, _ = a.train_iter(gt)
        except RuntimeError as e:
            assert "tile ranges" in str(e)
            hit = it
            break
        lb, _ = b.train_iter(gt)
        assert la == pytest.approx(lb, rel=1e-4), it
    if hit is None:
        pytest.skip("the Gaussians never crowded a tile beyond 256 in 200 iterations")
    assert not a.slots and a.step_index in (hit - 1, hit - 2)   # hit-2: step hit-1's own report already carried the flag
    while a.step_index < hit + 3:
        la, _ = a.train_iter(gt)
    while b.step_index < hit + 3:
        lb, _ = b.train_iter(gt)
    assert la == pytest.approx(lb, rel=1e-3)


def test_trajectory_follows_the_oracle():
    H, W, n, K = 256, 256, 10_000, 25
    init, gt, fit = _setup(H, W, n, seed=5, slots=True)
    o = _oracle_fit(H, W, init, gt[0])
    gtd = gt.to(cuda)
    losses, o_losses = [], []
    for _ in range(K):
        l, _psnr = fit.train_iter(gtd)
        losses.append(l)
        o_losses.append(o.step())
    assert losses[-1] < losses[0]
    np.testing.assert_allclose(losses, o_losses, rtol=5e-4)
    # parameters drift apart only through fp32 reduction order amplified by Adan's sign-like step
    for mine, ref in ((fit._xyz, o.p[0]), (fit._cholesky, o.p[1]), (fit._features_dc, o.p[2])):
        d = np.abs(mine.cpu().numpy() - ref)
        assert np.median(d) <= 1e-6 and (d > 1e-3).mean() <= 0.01


def test_uint8_target_equals_float_target():
    H, W, n = 128, 160, 2_000
    init, gt, fit_a = _setup(H, W, n, seed=9)
    _, _, fit_b = _setup(H, W, n, seed=9)
    u8 = (gt[0] * 255).round().to(torch.uint8)              # [3,H,W]
    gt_f = (u8.float() / 255).unsqueeze(0).contiguous()
    ga = torch.empty((n, 8), device=cuda)
    gb = torch.empty((n, 8), device=cuda)
    fit_a.train_iters(gt_f.to(cuda), 1, grads_out=ga)
    fit_b.train_iters(u8.permute(1, 2, 0).contiguous().to(cuda), 1, grads_out=gb)
    assert fit_a.loss() == pytest.approx(fit_b.loss(), rel=1e-6)
    assert _rel(ga.cpu().numpy(), gb.cpu().numpy()) <= 1e-5


def test_more_than_256_per_tile_truncates_like_the_reference():
    """trained-like Gaussians on a small frame: many tiles exceed 256 entries"""
    H, W, n = 96, 128, 6_000
    init, gt, fit = _setup(H, W, n, seed=11, trained_like=True)
    o = _oracle_fit(H, W, init, gt[0])
    o_loss, o_g, r = o.grads()
    tiles = fit.layout.tiles
    cnt = (r["tile_bins"][:tiles, 1] - r["tile_bins"][:tiles, 0])
    assert cnt.max() > 256 + 64, cnt.max()
    grads = torch.empty((n, 8), device=cuda)
    ids = torch.full((fit.cap,), -7, dtype=torch.int32, device=cuda)
    fit.train_iters(gt.to(cuda), 1, grads_out=grads, ids_out=ids)
    assert fit.status()[1] == 0
    I = r["num_intersects"]
    got = ids[:I].cpu().numpy()
    exp = r["gaussian_ids_sorted"].copy()
    for t in range(tiles):                                   # entries past the first 256 are never rendered
        lo, hi = r["tile_bins"][t]
        if hi - lo > 256:
            exp[lo + 256:hi] = -1
    assert np.array_equal(got, exp)
    assert abs(fit.loss() - o_loss) <= 5e-6 * o_loss
    g = grads.cpu().numpy()
    r_g, _ = _ref_gpu_param_grads(init, gt, H, W)
    for mine, ref, ora in ((g[:, 0:2], r_g[0], o_g[0]), (g[:, 2:5], r_g[1], o_g[1]), (g[:, 5:8], r_g[2], o_g[2])):
        assert _rel(mine, ref) <= GRAD_RTOL
        assert _rel(mine, ora) <= ORACLE_GRAD_RTOL


@pytest.mark.parametrize("H,W,n,trained", [(256, 256, 10_000, False), (1080, 1920, 50_000, False),
                                           (1080, 1920, 50_000, True), (96, 128, 6_000, True), (37, 53, 500, False)])
def test_render_matches_reference_kernels_and_oracle(H, W, n, trained):
    from gsvc_b200 import _lib
    from oracle import oracle as O
    from tests import ref_gpu
    init, gt, fit = _setup(H, W, n, seed=13, trained_like=trained, cap=(1 << 21) if trained else None)
    means = torch.from_numpy(_device_tanh(init["_xyz"].numpy()))
    L = init["_cholesky"] + torch.tensor([0.5, 0.0, 0.5])
    md, Ld, cd = means.to(cuda).contiguous(), L.to(cuda).contiguous(), init["_features_dc"].to(cuda).contiguous()
    rr = ref_gpu.pipeline(md, Ld, cd, torch.ones(n, 1, device=cuda), H, W)
    # through the module (raw parameters, clamped CHW image)
    img = fit.forward()[0].permute(1, 2, 0)
    assert (img - rr["out_img"].clamp(0, 1)).abs().max().item() <= 1e-5
    # through the C ABI on activated inputs: unclamped HWC image + final_idx + rendering order
    lay = fit.layout
    work = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=cuda)
    lib = _lib.load()
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.gs2d_fit_init(n, fit.cap, H, W, work.data_ptr(), lay.total_bytes, None, s))
    out = torch.empty((H, W, 3), device=cuda)
    fidx = torch.empty((H, W), dtype=torch.int32, device=cuda)
    ids = torch.empty((fit.cap,), dtype=torch.int32, device=cuda)
    _lib.check(lib.gs2d_render_chol(n, fit.cap, H, W, md.data_ptr(), Ld.data_ptr(), cd.data_ptr(), None, 0,
                                    work.data_ptr(), lay.total_bytes, out.data_ptr(), 0, 0, fidx.data_ptr(),
                                    ids.data_ptr(), s))
    torch.cuda.synchronize()
    I = rr["num_intersects"]
    st = work[lay.off_status:lay.off_status + 16].view(torch.int32).cpu()
    tiles = lay.tiles
    bins = rr["tile_bins"][:tiles].cpu().numpy()
    cnt = bins[:, 1] - bins[:, 0]
    assert (int(st[0]), int(st[1]), int(st[3])) == (I, 0, int(cnt.max()))
    off = work[lay.off_tile_off:lay.off_tile_off + 4 * (tiles + 1)].view(torch.int32).cpu().numpy()
    assert np.array_equal(off[1:] - off[:-1], cnt)                 # tile ranges == the reference's tile_bins
    assert np.array_equal(off[:-1][cnt > 0], bins[cnt > 0, 0])
    assert torch.equal(fidx, rr["final_idx"]), f"{(fidx != rr['final_idx']).sum().item()} final_idx pixels differ"
    assert (out - rr["out_img"]).abs().max().item() <= 1e-5
    exp = rr["gaussian_ids_sorted"].clone()
    for t in np.nonzero(bins[:, 1] - bins[:, 0] > 256)[0]:
        exp[bins[t, 0] + 256:bins[t, 1]] = -1
    assert torch.equal(ids[:I], exp)
    if H * W <= 256 * 256:   # independent CPU checker on the small configs (see the note above)
        r = O.render(means.numpy(), L.numpy(), init["_features_dc"].numpy(), np.ones((n, 1), np.float32), H, W)
        d = np.abs(out.cpu().numpy() - r["out_img"]).max(-1)
        flips = d > 1e-5 * max(1.0, float(np.abs(r["out_img"]).max()))
        assert flips.sum() <= 4
        assert (fidx.cpu().numpy() != r["final_idx"])[~flips].sum() == 0


def test_overflow_is_a_noop_and_recovers():
    H, W, n = 256, 256, 10_000
    init, gt, fit = _setup(H, W, n, seed=3, cap=4096)       # ~24k intersections needed
    gtd = gt.to(cuda)
    before = fit._xyz.clone()
    fit.train_iters(gtd, 3)
    I, overflow, _ = fit.status()
    assert overflow == 1 and I > 4096
    assert torch.equal(before, fit._xyz)
    with pytest.raises(RuntimeError, match="regrown"):
        fit.loss()
    assert fit.step_index == 0 and fit.cap == 8192
    for _ in range(3):                                       # 8192 -> 16384 -> 32768
        try:
            fit.train_iters(gtd, 2)
            fit.loss()
            break
        except RuntimeError:
            continue
    assert fit.status()[1] == 0 and fit.step_index == 2
    ref_init, _, ref = _setup(H, W, n, seed=3)
    ref.train_iters(gtd, 2)
    assert fit.loss() == pytest.approx(ref.loss(), rel=1e-5)


def test_state_dict_round_trip_with_reference_names():
    H, W, n = 64, 64, 500
    _, gt, fit = _setup(H, W, n)
    sd = fit.state_dict()
    assert set(sd) == {"_xyz", "_cholesky", "_features_dc"}
    assert sd["_xyz"].shape == (n, 2) and sd["_cholesky"].shape == (n, 3) and sd["_features_dc"].shape == (n, 3)


def test_in_place_parameter_edit_invalidates_the_precount():
    """every step pre-counts the next step's tiles; editing the parameters from Python in between
    must fall back to a fresh count (FusedFrameFit watches the tensors' version counters)"""
    H, W, n = 128, 160, 2_000
    init, gt, a = _setup(H, W, n, seed=21)
    _, _, b = _setup(H, W, n, seed=21)
    gtd = gt.to(cuda)
    a.train_iters(gtd, 2)
    b.train_iters(gtd, 2)
    shift = torch.zeros_like(a._xyz)
    shift[::2, 0] = 0.3                     # moves half of the Gaussians across tiles
    a._xyz.data.add_(shift)                 # .data edits do not bump the version: tell the module
    a._counted_for = None
    with torch.no_grad():
        b._xyz.add_(shift)                  # in-place edit: detected through ._version
    ia = torch.full((a.cap,), -7, dtype=torch.int32, device=cuda)
    ib = torch.full((b.cap,), -7, dtype=torch.int32, device=cuda)
    a.train_iters(gtd, 1, ids_out=ia)
    b.train_iters(gtd, 1, ids_out=ib)
    assert torch.equal(ia, ib)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-6)
    # and the order matches a fresh model holding the same parameters
    c = FusedFrameFitFrom(b, H, W, n)
    torch.cuda.synchronize()


def FusedFrameFitFrom(src, H, W, n):
    from gsvc_b200.fit import FusedFrameFit
    return FusedFrameFit(H, W, n, cuda, init={"_xyz": src._xyz.detach().cpu(), "_cholesky": src._cholesky.detach().cpu(),
                                              "_features_dc": src._features_dc.detach().cpu()})


def test_u8_target_all_values_match_ieee_division():
    """every uint8 value once per channel: the device's division-free value/255 must give the
    same loss as the float target holding IEEE value/255 (double accumulation, order-free to 1e-12)"""
    H, W, n = 16, 16, 64
    init, _, fa = _setup(H, W, n, seed=2)
    _, _, fb = _setup(H, W, n, seed=2)
    vals = torch.arange(256, dtype=torch.uint8).view(16, 16, 1).expand(16, 16, 3).contiguous()
    gt_f = (vals.float() / 255).permute(2, 0, 1).unsqueeze(0).contiguous()
    fa.train_iters(gt_f.to(cuda), 1)
    fb.train_iters(vals.to(cuda), 1)
    assert fa.loss() == pytest.approx(fb.loss(), rel=1e-12)


@pytest.mark.parametrize("fmt,lag,prefetch", [("rgb", 1, 2), ("rgb", 0, 2), ("i420", 1, 3), ("f32", 2, 4)])
def test_fit_stream_equals_resident_loop(fmt, lag, prefetch):
    """pinned-host frames streamed through the frame pipe (copies overlapped, results handed back
    `lag` steps late) give the same losses, in order, as the resident loop"""
    from gsvc_b200 import yuv
    H, W, n, K = 128, 160, 2_000, 13
    init, gt, a = _setup(H, W, n, seed=4)
    _, _, b = _setup(H, W, n, seed=4)
    u8 = (gt[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()
    rgbs = [((u8.int() + 3 * i) % 256).to(torch.uint8) for i in range(3)]                  # three different frames
    if fmt == "rgb":
        frames = [f.pin_memory() for f in rgbs]
    elif fmt == "i420":
        frames = [yuv.rgb_to_i420(f).contiguous().pin_memory() for f in rgbs]
    else:
        frames = [(f.permute(2, 0, 1).float() / 255).unsqueeze(0).contiguous().pin_memory() for f in rgbs]
    seq = [frames[i % 3] for i in range(K)]
    la = [l for l, _ in a.fit_stream(iter(seq), prefetch=prefetch, lag=lag)]
    assert len(la) == K
    with pytest.raises(ValueError):
        next(a.fit_stream([frames[0].clone()]))          # not pinned
    lb = []
    for f in seq:
        l, _ = b.train_iter(f.to(cuda))
        lb.append(l)
    np.testing.assert_allclose(la, lb, rtol=1e-5)
    assert a.step_index == K


@pytest.mark.parametrize("exchange", ["nccl", "peer", "sharded", "sharded-pull", "sharded-remote-scatter"])
@pytest.mark.parametrize("H,W,n,world", [(256, 256, 10_000, 2), (200, 328, 3_000, 3), (1080, 1920, 50_000, 4)])
def test_tile_row_bands_sum_to_the_whole_frame(H, W, n, world, exchange):
    """spatial mode on ONE GPU: every band's gradients (and loss) added into the same buffer in
    place of the all-reduce must reproduce the whole-frame step"""
    from gsvc_b200.fit import SpatialFrameFit
    from gsvc_b200 import synth
    init, gt, full = _setup(H, W, n, seed=17)
    # "nccl": all bands add into one buffer (what the all-reduce would produce); "peer": one buffer
    # per emulated rank, summed in rank order inside the update kernel (the NVLink exchange path);
    # "sharded": as "peer", but rank r updates only its share of the Gaussians and writes their next
    # records into the tile slots of the rank that owns each tile row (one workspace per rank)
    # (default: the update kernel scatters the next records into the owning rank's slots itself;
    # "-local-scatter": every rank re-projects all Gaussians for its band instead); bands balanced by work
    # "sharded" = pushed reduce-scatter (the tile kernels add into the owner's buffer); "-pull": the update kernel
    # reads every rank's rows instead; "-remote-scatter": it also writes the next records into the owners' slots
    banded = SpatialFrameFit(H, W, n, cuda, init=init, emulate_world=world, exchange=exchange.split("-")[0],
                             remote_scatter=exchange.endswith("remote-scatter"), push=not exchange.endswith("pull"),
                             balance=(world == 3))
    gtd = gt.to(cuda)
    ga = torch.empty((n, 8), device=cuda)
    gb = torch.empty((n, 8), device=cuda)
    full.train_iters(gtd, 1, grads_out=ga)
    banded.train_iters(gtd, 1, grads_out=gb)
    assert banded.loss() == pytest.approx(full.loss(), rel=2e-6)
    a, b = ga.cpu().numpy(), gb.cpu().numpy()
    for sl in (slice(0, 2), slice(2, 5), slice(5, 8)):
        assert _rel(b[:, sl], a[:, sl]) <= GRAD_RTOL
    for _ in range(4):
        full.train_iters(gtd, 1)
        banded.train_iters(gtd, 1)
    assert banded.loss() == pytest.approx(full.loss(), rel=1e-4)
    d = (banded._features_dc - full._features_dc).abs()
    assert float(d.median()) <= 1e-6


def _crowding_setup(slots=True, lr=0.05):
    """Gaussians that drift into one tile until it holds more than 256 (the recipe of
    test_slot_overflow_in_a_later_step)"""
    H, W, n = 64, 64, 3_000
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit
    init = synth.gaussians(n, seed=2)
    init["_xyz"] = init["_xyz"] * 0.05
    init["_cholesky"] = init["_cholesky"] * 0 + torch.tensor([1.2, 0.0, 1.2])
    gt = synth.target_frame(H, W, seed=2)
    return H, W, n, init, gt, FusedFrameFit(H, W, n, cuda, init=init, slots=slots, lr=lr)


def test_train_until_absorbs_a_slot_overflow():
    """the recovery is part of the API -- every requested iteration really happens, with
    the results of a model that ran on tile ranges from the start"""
    from gsvc_b200.fit import GS2DOverflow
    H, W, n, init, gt, a = _crowding_setup(slots=True)
    _, _, _, _, _, b = _crowding_setup(slots=False)
    gtd = gt.to(cuda)
    K = 12       # (short: a tile holding every Gaussian at lr 0.05 amplifies rounding noise quickly)
    la = a.train_until(gtd, K)
    b.train_iters(gtd, K)
    lb = b.loss()
    assert a.step_index == K and b.step_index == K
    assert not a.slots, "the recipe crowds a tile beyond 256 at once"
    assert la == pytest.approx(lb, rel=1e-3)
    assert issubclass(GS2DOverflow, RuntimeError)
    # checking every few iterations gives the same thing; so does a long run, as far as the count goes
    _, _, _, _, _, c = _crowding_setup(slots=True)
    lc = c.train_until(gtd, K, check_every=5)
    assert c.step_index == K and lc == pytest.approx(lb, rel=1e-3)
    c.train_until(gtd, 150)
    assert c.step_index == 150 and math.isfinite(c.loss())


def test_fit_stream_resumes_after_an_overflow():
    """frames of the steps that became no-ops are issued again, none is lost, none twice"""
    H, W, n, init, gt, a = _crowding_setup(slots=True)
    _, _, _, _, _, b = _crowding_setup(slots=False)
    host = gt.contiguous().pin_memory()
    K = 40
    la = [l for l, _ in a.fit_stream(host for _ in range(K))]
    lb = [l for l, _ in b.fit_stream(host for _ in range(K))]
    assert len(la) == K and len(lb) == K and a.step_index == K and not a.slots
    np.testing.assert_allclose(la[:12], lb[:12], rtol=1e-3)
    assert la[-1] < la[0] and all(math.isfinite(x) for x in la)


def test_gop_fit_survives_a_slot_overflow():
    """fit_gop / fit_sequence on a model that overflows a slot mid-fit: every frame
    receives all of its iterations"""
    from gsvc_b200 import gop
    from gsvc_b200.fit import FusedFrameFit
    H, W, n, init, gt, _ = _crowding_setup()
    gtd = gt.to(cuda)
    made = []

    def make(prev):
        f = FusedFrameFit(H, W, n, cuda, init=init if prev is None else prev, slots=True, lr=0.05)
        made.append(f)
        return f

    out = gop.fit_sequence(lambda f: gtd, 3, [0], make, 120, 40, seed=None)
    assert sorted(out) == [0, 1, 2]
    assert [out[f]["iters"] for f in range(3)] == [120, 40, 40]
    assert [m.step_index for m in made] == [120, 40, 40]
    assert all(math.isfinite(out[f]["psnr"]) for f in out)


def test_step_lr_matches_torch_scheduler_across_a_boundary():
    """iteration `it` runs at lr * gamma^((it-1) // step_size), as optimizer.step() followed by
    StepLR.step() gives (GaussianSplats_Represent.py:52, 203-206)"""
    from gsvc_b200.fit import FusedFrameFit
    fit = FusedFrameFit(32, 32, 64, cuda, lr=1e-2, lr_step_size=5, lr_gamma=0.5)
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1e-2)
    sch = torch.optim.lr_scheduler.StepLR(opt, step_size=5, gamma=0.5)
    gt = torch.zeros(1, 3, 32, 32, device=cuda)
    for it in range(1, 14):
        used = opt.param_groups[0]["lr"]       # what optimizer.step() of iteration `it` uses
        fit.step_index = it                    # as _issue has it while the step is being issued
        assert fit.current_lr() == pytest.approx(used, rel=1e-12), it
        opt.step()
        sch.step()
    fit.step_index = 0
    fit.train_iters(gt, 7)
    assert fit.current_lr() == pytest.approx(1e-2 * 0.5, rel=1e-12)    # iteration 7 -> (7-1)//5 = 1


def test_parameters_are_column_views_of_one_packed_buffer():
    """SURVEY 8 a12 / north_star "16-byte loads": the master copy is fp32 [n][8] (xyz 2 | chol 3 | feat 3);
    the reference's names and shapes are views of it, and a tensor assigned from outside is packed again"""
    H, W, n = 96, 128, 3_000
    init, gt, a = _setup(H, W, n, seed=4, slots=True)
    _, _, b = _setup(H, W, n, seed=4, slots=True)
    p8 = a._params8
    assert p8.shape == (n, 8) and p8.is_contiguous() and p8.data_ptr() % 16 == 0
    assert a._xyz.shape == (n, 2) and a._xyz.stride() == (8, 1) and a._xyz.data_ptr() == p8.data_ptr()
    assert a._cholesky.data_ptr() == p8.data_ptr() + 8 and a._features_dc.data_ptr() == p8.data_ptr() + 20
    assert torch.equal(p8[:, 2:5], init["_cholesky"].to(cuda))
    gtd = gt.to(cuda)
    a.train_iters(gtd, 3)
    b.train_iters(gtd, 3)
    d = (a._params8 - b._params8).abs()       # two runs differ by the order of the gradient atomics only
    assert float(d.median()) <= 1e-6 and float((d > 1e-4).float().mean()) <= 0.01
    sd = a.state_dict()
    assert all(sd[k].is_contiguous() for k in ("_xyz", "_cholesky", "_features_dc"))
    # somebody swaps a parameter tensor instead of writing into it
    new_xyz = a._xyz.detach().clone() * 0.9
    a._xyz = torch.nn.Parameter(new_xyz, requires_grad=False)
    with torch.no_grad():
        b._xyz.copy_(new_xyz)
    a.train_iters(gtd, 1)
    b.train_iters(gtd, 1)
    assert a._xyz.data_ptr() == a._params8.data_ptr()
    assert a.loss() == pytest.approx(b.loss(), rel=1e-6)


@pytest.mark.parametrize("fmt", ["f32", "u8", "i420"])
def test_prepared_target_gives_the_same_bits_as_the_in_kernel_decode(fmt):
    """gs2d_fit_target_prepare (the frame converted once into the tile kernel's order) against the
    per-iteration decode of the same frame: identical loss sums and gradients"""
    from gsvc_b200 import yuv
    H, W, n = 200, 328, 3_000
    init, gt, a = _setup(H, W, n, seed=6, slots=True)
    _, _, b = _setup(H, W, n, seed=6, slots=True)
    b.prepare_targets = False
    rgb8 = (gt[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()
    frame = {"f32": gt, "u8": rgb8, "i420": yuv.rgb_to_i420(rgb8).contiguous()}[fmt].to(cuda)
    ga = torch.empty((n, 8), device=cuda)
    gb = torch.empty((n, 8), device=cuda)
    a.train_iters(frame, 1, grads_out=ga)
    b.train_iters(frame, 1, grads_out=gb)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-12)     # same per-tile sums, added up in another order
    assert (ga - gb).abs().max().item() <= 1e-4 * gb.abs().max().item()
    # an in-place edit of the frame is seen (version counter), a new tensor too
    frame2 = frame.clone()
    if fmt == "f32":
        frame2.mul_(0.5)
    else:
        frame2.div_(2, rounding_mode="floor")
    a.train_iters(frame2, 1)
    b.train_iters(frame2, 1)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-6)
    with torch.no_grad():
        frame2.zero_()
    a.train_iters(frame2, 1)
    b.train_iters(frame2, 1)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-6)


def test_intersections_per_tile_row_match_the_binning():
    """the weights the tile-row split is balanced with are the per-row sums of the tile counts"""
    _need = pytest.importorskip("torch")
    from tests import ref_gpu
    from gsvc_b200.fit import _intersections_per_tile_row, split_tile_rows
    import gsvc_b200.cuda as C
    H, W, n = 520, 392, 7_000
    from gsvc_b200 import synth
    g = synth.gaussians(n, seed=9, trained_like=True)
    rows = _intersections_per_tile_row(g["_xyz"].to(cuda), g["_cholesky"].to(cuda), H, W)
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    xys, depths, radii, conics, nth = C.project_gaussians_2d_forward(n, g["means2d"].to(cuda), g["L"].to(cuda), H, W, tb, 0.01)
    cum = torch.cumsum(nth, 0, dtype=torch.int32)
    I = int(cum[-1])
    isect, _ids = C.map_gaussian_to_intersects(n, I, xys, depths, radii, cum, tb)
    tile_row = ((isect >> 32) // tb[0]).long()
    want = torch.bincount(tile_row, minlength=tb[1]).tolist()
    assert [int(r) for r in rows] == want
    bands = split_tile_rows(tb[1], 4, weights=rows)
    load = [sum(rows[a:b]) for a, b in bands]
    assert max(load) <= 1.35 * (sum(rows) / 4)


def test_fused_adan_arithmetic_element_by_element_over_several_steps():
    """VERDICT r1 (weak #2): at step 1 Adan moves every element by +-lr, which hides the arithmetic behind
    the sign of near-zero gradients.  Here the update is checked in isolation from step 1 to step 5: the
    gradients the fused step itself handed to its optimizer (grads_out) drive a float32 restatement of the
    reference's update (optimizer.py:296-362: exp_avg, exp_avg_diff, exp_avg_sq, bias corrections, eps outside
    the square root) from the same start; parameters must agree element by element to a few ulps of lr."""
    H, W, n = 256, 256, 10_000
    init, gt, fit = _setup(H, W, n, seed=5, slots=True)
    gtd = gt.to(cuda)
    p = np.concatenate([init[k].numpy() for k in ("_xyz", "_cholesky", "_features_dc")], axis=1).astype(np.float32)
    m = np.zeros_like(p); v = np.zeros_like(p); d = np.zeros_like(p); prev = None
    f = np.float32
    b1, b2, b3, lr, eps = 0.98, 0.92, 0.99, 1e-3, 1e-8
    g_dev = torch.empty((n, 8), device=cuda)
    worst = 0.0
    for t in range(1, 6):
        fit.train_iters(gtd, 1, grads_out=g_dev)
        g = g_dev.cpu().numpy()
        if prev is None:
            prev = g.copy()
        diff = g - prev
        m = (f(b1) * m + f(1 - b1) * g).astype(f)
        d = (f(b2) * d + f(1 - b2) * diff).astype(f)
        u = g + f(b2) * diff
        v = (f(b3) * v + f(1 - b3) * u * u).astype(f)
        denom = np.sqrt(v) / f(math.sqrt(1 - b3 ** t)) + f(eps)
        p = (p - f(lr / (1 - b1 ** t)) * m / denom - f(lr * b2 / (1 - b2 ** t)) * d / denom).astype(f)
        prev = g
        got = fit._params8.cpu().numpy()
        err = np.abs(got - p)
        worst = max(worst, float(err.max()))
        # the device multiplies by 1/sqrt(1-b3^t) where this divides, and contracts a*b+c: a few ulps of the
        # update (|update| <= ~lr), which after the final subtraction is at most one unit in the last place
        # of the parameter itself (2.4e-7 for |p| in [2, 4)) -- never more
        ulp = np.spacing(np.maximum(np.abs(p), np.abs(got)).astype(f))
        assert err.max() <= 1e-6, (t, float(err.max()))                    # 0.1 % of lr at the very worst
        assert (err > ulp + 1e-9).mean() <= 0.01, (t, float((err > ulp + 1e-9).mean()), float(err.max()))
        p = got.copy()      # re-anchor: the next step starts from the device's parameters (isolates each step)
        m_dev = fit.adan_state[0].cpu().numpy(); v_dev = fit.adan_state[1].cpu().numpy(); d_dev = fit.adan_state[2].cpu().numpy()
        assert np.abs(m_dev - m).max() <= 1e-6 * max(1.0, np.abs(m).max())
        assert np.abs(d_dev - d).max() <= 1e-6 * max(1.0, np.abs(d).max())
        assert np.abs(v_dev - v).max() <= 1e-6 * max(1e-12, np.abs(v).max())
        m, v, d = m_dev, v_dev, d_dev
