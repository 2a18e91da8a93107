"""Decode side of the codec on the GPU: a frame's Gaussians in the quantised wire format
(GaussianSplats_Compress.py:34-37, 71-84 of the reference: fp16 means, 6-bit Cholesky codes with a
per-channel scale / beta, two 3-bit residual-VQ colour indices into two 8-entry codebooks) rendered
straight from the codes -- 9 bytes per Gaussian read by the projection, no fp32 copy of the frame.

    q = QuantizedFrame.from_state(state, codebooks)     # test / synthetic data: the encode side
    img = q.render(H, W)                                # [1,3,H,W] clamped, as forward_quantize's render
    state = q.dequantize()                              # _xyz, _cholesky, _features_dc a P-frame fit starts from

Entropy (de)coding of the code streams and the training of the codebooks are the reference's
third-party dependencies (constriction, vector_quantize_pytorch) and stay where they are."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch


@dataclass
class QuantizedFrame:
    xyz_half: torch.Tensor      # float16 [n,2] (CUDA)
    chol_code: torch.Tensor     # uint8 [n,3], 0..63
    chol_scale: torch.Tensor    # fp32 [3] (host or device)
    chol_beta: torch.Tensor     # fp32 [3]
    color_index: torch.Tensor   # uint8 [n,2], 0..7
    codebooks: torch.Tensor     # fp32 [2,8,3] (CUDA)

    def __post_init__(self):
        dev = self.xyz_half.device
        if dev.type != "cuda":
            raise ValueError("the quantised frame must live on a CUDA device")
        n = self.xyz_half.shape[0]
        if self.xyz_half.dtype != torch.float16 or self.xyz_half.shape != (n, 2):
            raise ValueError("xyz_half must be float16 [n,2]")
        if self.chol_code.dtype != torch.uint8 or self.chol_code.shape != (n, 3):
            raise ValueError("chol_code must be uint8 [n,3]")
        if self.color_index.dtype != torch.uint8 or self.color_index.shape != (n, 2):
            raise ValueError("color_index must be uint8 [n,2]")
        if self.codebooks.dtype != torch.float32 or self.codebooks.shape != (2, 8, 3):
            raise ValueError("codebooks must be fp32 [2,8,3]")
        self.xyz_half, self.chol_code = self.xyz_half.contiguous(), self.chol_code.contiguous()
        self.color_index, self.codebooks = self.color_index.contiguous(), self.codebooks.contiguous().to(dev)
        self._scale = [float(v) for v in self.chol_scale.detach().cpu().reshape(3)]
        self._beta = [float(v) for v in self.chol_beta.detach().cpu().reshape(3)]
        self._work = None

    @property
    def n(self) -> int:
        return int(self.xyz_half.shape[0])

    @property
    def wire_bytes(self) -> int:
        """bytes the projection reads per frame (codes + side information; before entropy coding)"""
        return 9 * self.n + 24 + 192

    def _struct(self):
        from . import _lib
        q = _lib.QuantizedFrame()
        q.xyz_half, q.chol_code = self.xyz_half.data_ptr(), self.chol_code.data_ptr()
        q.color_index, q.codebooks = self.color_index.data_ptr(), self.codebooks.data_ptr()
        for c in range(3):
            q.chol_scale[c], q.chol_beta[c] = self._scale[c], self._beta[c]
        return q

    @classmethod
    def from_state(cls, state: dict, codebooks: torch.Tensor, device=None) -> "QuantizedFrame":
        """the encode side, for synthetic / test data: fp16 cast, UniformQuantizer._init_data +
        compress (quantize.py:41-47, 82-84), and nearest-entry residual VQ of the colours against
        the GIVEN codebooks (what ResidualVQ does at eval time with trained codebooks)"""
        device = torch.device(device or state["_xyz"].device)
        xyz = state["_xyz"].detach().float().to(device)
        chol = state["_cholesky"].detach().float().to(device)
        feat = state["_features_dc"].detach().float().to(device)
        books = codebooks.detach().float().to(device)
        beta = chol.min(dim=0)[0]
        scale = (chol.max(dim=0)[0] - beta) / 63
        code = ((chol - beta) / scale).clamp(0, 63).round().to(torch.uint8)
        i0 = torch.cdist(feat, books[0]).argmin(dim=1)
        i1 = torch.cdist(feat - books[0][i0], books[1]).argmin(dim=1)
        return cls(xyz.half(), code, scale, beta, torch.stack([i0, i1], dim=1).to(torch.uint8), books)

    def _workspace(self, H: int, W: int, cap: int):
        from . import _lib
        from . import cuda as _C
        key = (H, W, cap)
        if self._work is None or self._work[0] != key:
            lay = _lib.FitLayout()
            lib = _lib.load()
            _lib.check(lib.gs2d_fit_layout_for(self.n, cap, H, W, lay), "fit_layout_for")
            work = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=self.xyz_half.device)
            with _C._guard(work.device):
                _lib.check(lib.gs2d_fit_init(self.n, cap, H, W, work.data_ptr(), lay.total_bytes, None,
                                             torch.cuda.current_stream(work.device).cuda_stream), "fit_init")
            self._work = (key, work)
        return self._work[1]

    @torch.no_grad()
    def render(self, H: int, W: int, cap: int | None = None, final_idx: torch.Tensor | None = None) -> torch.Tensor:
        """[1,3,H,W] fp32 clamped to [0,1] (forward_quantize's "render", GaussianSplats_Compress.py:80-83)"""
        from . import _lib
        from . import cuda as _C
        dev = self.xyz_half.device
        cap = int(cap) if cap else max(8 * self.n, 1 << 18)
        work = self._workspace(H, W, cap)
        out = torch.empty((1, 3, H, W), dtype=torch.float32, device=dev)
        q = self._struct()
        with _C._guard(dev):
            _lib.check(_lib.load().gs2d_render_quantized(
                self.n, cap, H, W, ctypes.addressof(q), work.data_ptr(), work.numel(), out.data_ptr(), 1, 1,
                None if final_idx is None else final_idx.data_ptr(), None,
                torch.cuda.current_stream(dev).cuda_stream), "render_quantized")
        return out

    @torch.no_grad()
    def dequantize(self) -> dict:
        """{_xyz, _cholesky, _features_dc}: the fp32 parameters the codes stand for"""
        from . import _lib
        from . import cuda as _C
        dev, n = self.xyz_half.device, self.n
        out = {"_xyz": torch.empty((n, 2), device=dev), "_cholesky": torch.empty((n, 3), device=dev),
               "_features_dc": torch.empty((n, 3), device=dev)}
        q = self._struct()
        with _C._guard(dev):
            _lib.check(_lib.load().gs2d_dequantize(n, ctypes.addressof(q), out["_xyz"].data_ptr(),
                                                   out["_cholesky"].data_ptr(), out["_features_dc"].data_ptr(),
                                                   torch.cuda.current_stream(dev).cuda_stream), "dequantize")
        return out
