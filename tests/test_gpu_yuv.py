"""GPU parity of the I420 frame path: gs2d_i420_to_rgb and the fused fit step's I420 target
format against the OpenCV-pinned restatement (oracle/yuv.py); bit-exact."""
import hashlib
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda:0")
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "yuv.npz"))


def test_converter_on_all_triples_and_golden():
    from gsvc_b200 import yuv
    from oracle import yuv as Y
    frame, H, W = Y.all_triples_frame()
    rgb = yuv.i420_to_rgb(torch.from_numpy(frame).to(cuda), H, W).cpu().numpy()
    assert hashlib.sha256(np.ascontiguousarray(rgb).tobytes()).hexdigest() == str(G["all_triples_sha256"])
    small = yuv.i420_to_rgb(torch.from_numpy(G["small_yuv"]).to(cuda), 48, 64).cpu().numpy()
    assert np.array_equal(small, G["small_rgb"])
    with pytest.raises(ValueError):
        yuv.i420_to_rgb(torch.zeros(18, dtype=torch.uint8, device=cuda), 3, 4)


@pytest.mark.parametrize("H,W,n", [(256, 256, 8000), (1080, 1920, 50_000), (130, 202, 2000)])
def test_fit_step_on_the_raw_frame_equals_fit_step_on_its_rgb(H, W, n):
    """same model, same frame: once as raw I420 (decoded in the tile kernel), once as the RGB frame
    the oracle decodes from it -> identical loss and identical gradients handed to Adan"""
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit
    from oracle import yuv as Y
    rng = np.random.default_rng(H + W)
    raw = rng.integers(0, 256, size=H * W * 3 // 2, dtype=np.uint8)
    rgb = torch.from_numpy(Y.i420_to_rgb(raw, H, W)).to(cuda)
    init = synth.gaussians(n, seed=9)
    a = FusedFrameFit(H, W, n, cuda, init=init)
    b = FusedFrameFit(H, W, n, cuda, init=init)
    ga, gb = torch.empty((n, 8), device=cuda), torch.empty((n, 8), device=cuda)
    a.train_iters(torch.from_numpy(raw).view(H * 3 // 2, W).to(cuda), 1, grads_out=ga)
    b.train_iters(rgb, 1, grads_out=gb)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-12)         # same fp32 terms, double accumulation
    d = (ga - gb).abs().max() / gb.abs().max()
    assert float(d) <= 1e-5                                       # atomic order only
    for _ in range(3):
        a.train_iters(torch.from_numpy(raw).to(cuda), 1)          # flat works too
        b.train_iters(rgb, 1)
    assert a.loss() == pytest.approx(b.loss(), rel=1e-5)


def test_process_yuv_video_drop_in(tmp_path):
    from gsvc_b200 import yuv
    from oracle import yuv as Y
    rng = np.random.default_rng(2)
    H, W = 36, 64
    frames = [rng.integers(0, 256, size=H * W * 3 // 2, dtype=np.uint8) for _ in range(3)]
    p = tmp_path / "clip_64x36.yuv"
    with open(p, "wb") as f:
        for fr in frames:
            f.write(fr.tobytes())
        f.write(b"\\x00" * 100)                  # trailing partial frame: ignored, as the reference
    out = yuv.process_yuv_video(str(p), W, H)
    assert len(out) == 3
    for got, fr in zip(out, frames):
        assert got.dtype == np.uint8 and np.array_equal(got, Y.i420_to_rgb(fr, H, W))
