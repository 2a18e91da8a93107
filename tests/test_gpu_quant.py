"""GPU parity of the decode side: gs2d_dequantize / gs2d_render_quantized against the
reference-pinned restatement (oracle/quant.py) and against the fp32 render of the same Gaussians."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda:0")
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "quant.npz"))


def _golden_frame():
    from gsvc_b200.decode import QuantizedFrame
    t = lambda k: torch.from_numpy(G[k])
    return QuantizedFrame(torch.from_numpy(G["xyz_half_bits"].view(np.float16).copy()).to(cuda), t("chol_code").to(cuda),
                          t("chol_scale"), t("chol_beta"), t("color_index").to(cuda), t("codebooks").to(cuda))


def test_dequantize_is_bit_exact_with_the_reference_classes():
    d = _golden_frame().dequantize()
    assert np.array_equal(d["_xyz"].cpu().numpy(), G["xyz_q"])
    assert np.array_equal(d["_cholesky"].cpu().numpy(), G["chol_deq"])
    assert np.array_equal(d["_features_dc"].cpu().numpy(), G["colours"])


@pytest.mark.parametrize("H,W,n", [(256, 256, 4096), (1080, 1920, 50_000), (130, 202, 1500)])
def test_render_from_codes_equals_render_from_the_dequantised_tensors(H, W, n):
    """same Gaussians, once read as 9-byte codes inside the projection, once as the fp32 tensors the
    codes stand for (the reference's forward_quantize feeds those to the two operators): the image
    and the per-pixel last index must be identical"""
    from gsvc_b200 import synth
    from gsvc_b200.decode import QuantizedFrame
    from gsvc_b200.fit import FusedFrameFit
    from oracle import quant as Q
    init = synth.gaussians(n, seed=4)
    books = torch.rand(2, 8, 3, generator=torch.Generator().manual_seed(1)) * torch.tensor([1.0, 0.3]).view(2, 1, 1)
    q = QuantizedFrame.from_state(init, books, device=cuda)
    assert q.wire_bytes < 0.3 * 32 * n
    fa = torch.empty((H, W), dtype=torch.int32, device=cuda)
    img = q.render(H, W, final_idx=fa)
    # the restatement's view of the same codes
    want = {"_xyz": Q.dequantize_xyz(q.xyz_half.cpu().numpy().view(np.uint16)),
            "_cholesky": Q.dequantize_cholesky(q.chol_code.cpu().numpy(), q._scale, q._beta),
            "_features_dc": Q.decode_colours(q.color_index.cpu().numpy(), books.numpy())}
    d = q.dequantize()
    for k in want:
        assert np.array_equal(d[k].cpu().numpy(), want[k]), k
    ref = FusedFrameFit(H, W, n, cuda, init={k: torch.from_numpy(v) for k, v in want.items()})
    assert torch.equal(img, ref.forward())
    # and it is a faithful picture of the unquantised model (6-bit Cholesky, 64 colours): loose
    full = FusedFrameFit(H, W, n, cuda, init=init).forward()
    assert float((img - full).abs().mean()) < 0.2


def test_bad_frames_are_rejected():
    from gsvc_b200.decode import QuantizedFrame
    z = torch.zeros
    with pytest.raises(ValueError):
        QuantizedFrame(z(4, 2), z(4, 3, dtype=torch.uint8), z(3), z(3), z(4, 2, dtype=torch.uint8), z(2, 8, 3))   # CPU
    with pytest.raises(ValueError):
        QuantizedFrame(z(4, 2, device=cuda), z(4, 3, dtype=torch.uint8, device=cuda), z(3), z(3),
                       z(4, 2, dtype=torch.uint8, device=cuda), z(2, 8, 3, device=cuda))                        # not fp16
