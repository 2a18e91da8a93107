"""Golden vectors for the decode side (wire format -> Gaussians) from the reference's OWN quantizer
code (quantize.py), run on CPU in the build container:  python tests/golden/make_golden_quant.py
quantize.py imports three packages that are not installed here (vector_quantize_pytorch,
constriction, and the reference's utils -> cv2 / pytorch_msssim); they are stubbed -- the three
pieces of arithmetic that define the decode live in quantize.py itself and run unmodified:
  FakeQuantizationHalf.forward            quantize.py:15-20    x.half().float()
  UniformQuantizer._init_data / compress / decompress   :41-47, :82-88   code*scale + beta
  VectorQuantizer.decompress              quantize.py:146-150  0 + e0[i0] + e1[i1]
(the residual-VQ ENCODER and its codebook training are the third-party package's and are not
restated: the fixture uses random codebooks and indices) -> quant.npz"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _stubs():
    v = types.ModuleType("vector_quantize_pytorch")
    v.VectorQuantize = v.ResidualVQ = object
    c = types.ModuleType("constriction")
    u = types.ModuleType("utils")
    u.__all__ = []
    sys.modules.update({"vector_quantize_pytorch": v, "constriction": c, "utils": u})
    sys.path.insert(0, "/root/reference")


if __name__ == "__main__":
    _stubs()
    import quantize as Q
    torch.manual_seed(11)
    n = 4096
    xyz = torch.atanh(2 * (torch.rand(n, 2) - 0.5))
    xyz[0] = torch.tensor([1e-8, -65504.0 * 0 + 3.0])          # a denormal-in-fp16 and a plain value
    xyz_q = Q.FakeQuantizationHalf.apply(xyz)
    chol = torch.rand(n, 3) * torch.tensor([3.0, 1.5, 2.0]) - torch.tensor([0.0, 0.7, 0.0])
    uq = Q.UniformQuantizer(signed=False, bits=6, learned=True, num_channels=3)
    uq._init_data(chol)
    code, deq = uq.compress(chol)
    assert torch.equal(uq.decompress(code.detach()), deq.detach())
    books = torch.rand(2, 8, 3) * torch.tensor([1.0, 0.25]).view(2, 1, 1)
    idx = torch.randint(0, 8, (n, 2))
    layers = [types.SimpleNamespace(_codebook=types.SimpleNamespace(embed=books[i:i + 1])) for i in range(2)]
    fake = types.SimpleNamespace(quantizer=types.SimpleNamespace(layers=layers))
    colours = Q.VectorQuantizer.decompress(fake, idx)
    np.savez_compressed(
        os.path.join(HERE, "quant.npz"),
        xyz=xyz.numpy(), xyz_q=xyz_q.numpy(), xyz_half_bits=xyz.half().view(torch.int16).numpy().view(np.uint16),
        chol=chol.numpy(), chol_code=code.detach().numpy().astype(np.uint8), chol_deq=deq.detach().numpy(),
        chol_scale=uq.scale.detach().numpy(), chol_beta=uq.beta.detach().numpy(),
        codebooks=books.numpy(), color_index=idx.numpy().astype(np.uint8), colours=colours.numpy())
    print("codes", int(code.min()), "..", int(code.max()), "| colours", tuple(colours.shape), "| wrote quant.npz")
