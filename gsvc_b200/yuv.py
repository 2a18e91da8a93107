"""Frame loading for the fit loop: raw planar I420 (the reference's .yuv input) on the GPU.

The reference decodes every frame on the CPU with OpenCV (`process_yuv_video`, utils.py:134-156:
np.frombuffer -> cv2.cvtColor(COLOR_YUV2RGB_I420)) and later turns the RGB array into an fp32
tensor.  Here the raw frame goes to the device as it is (1.5 bytes per pixel instead of 3, or 12
as fp32) and is either consumed directly by the fused fit step (target format I420: conversion
inside the tile kernel) or converted by gs2d_i420_to_rgb for callers that want the RGB frame.
Both are bit-identical to cv2 (tests/test_yuv_cpu.py, tests/test_gpu_yuv.py)."""
from __future__ import annotations

import os
from typing import Iterator, List

import numpy as np
import torch


def frame_bytes(width: int, height: int) -> int:
    return width * height * 3 // 2      # utils.py:136


def read_i420_frames(file_path: str, width: int, height: int, pin: bool = True) -> Iterator[torch.Tensor]:
    """the raw frames of a .yuv file, each a uint8 tensor [height*3/2, width] (the array the
    reference hands to cv2, utils.py:148) in pinned host memory"""
    size = frame_bytes(width, height)
    total = os.path.getsize(file_path) // size          # :138-139: a trailing partial frame is dropped
    with open(file_path, "rb") as f:
        for _ in range(total):
            buf = f.read(size)
            if len(buf) < size:
                break
            t = torch.from_numpy(np.frombuffer(buf, dtype=np.uint8).copy()).view(height * 3 // 2, width)
            yield t.pin_memory() if pin else t


def i420_to_rgb(yuv: torch.Tensor, height: int, width: int) -> torch.Tensor:
    """CUDA uint8 I420 frame -> uint8 [height, width, 3] RGB on the same device (cv2.COLOR_YUV2RGB_I420)"""
    from . import _lib
    from . import cuda as _C
    if yuv.dtype != torch.uint8 or not yuv.is_cuda or yuv.numel() != frame_bytes(width, height):
        raise ValueError("expected a CUDA uint8 I420 frame of width*height*3/2 bytes")
    if height % 2 or width % 2:
        raise ValueError("I420 needs even width and height")
    yuv = yuv.contiguous()
    rgb = torch.empty((height, width, 3), dtype=torch.uint8, device=yuv.device)
    with _C._guard(yuv.device):
        _lib.check(_lib.load().gs2d_i420_to_rgb(height, width, yuv.data_ptr(), rgb.data_ptr(),
                                                torch.cuda.current_stream(yuv.device).cuda_stream), "i420_to_rgb")
    return rgb


def process_yuv_video(file_path: str, width: int, height: int, device="cuda") -> List[np.ndarray]:
    """drop-in for the reference's process_yuv_video (utils.py:134-156): list of RGB uint8
    [height, width, 3] numpy frames, decoded on the GPU"""
    out = []
    for raw in read_i420_frames(file_path, width, height):
        out.append(i420_to_rgb(raw.to(device, non_blocking=True), height, width).cpu().numpy())
    return out


def rgb_to_i420(rgb: torch.Tensor) -> torch.Tensor:
    """synthetic data only: an I420 frame whose decode is close to `rgb` (uint8 [H,W,3]); BT.601
    limited range, chroma averaged over 2x2.  Not part of the reference's path (it only decodes)."""
    H, W, _ = rgb.shape
    r, g, b = (rgb[..., i].float() for i in range(3))
    y = 16 + 0.257 * r + 0.504 * g + 0.098 * b
    u = 128 - 0.148 * r - 0.291 * g + 0.439 * b
    v = 128 + 0.439 * r - 0.368 * g - 0.071 * b
    pool = lambda p: p.view(H // 2, 2, W // 2, 2).mean(dim=(1, 3))
    q = lambda p: p.round().clamp(0, 255).to(torch.uint8).reshape(-1)
    return torch.cat([q(y), q(pool(u)), q(pool(v))]).view(H * 3 // 2, W)
