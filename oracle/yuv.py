"""CPU restatement of the reference's frame loader arithmetic: planar I420 (YUV 4:2:0, 8 bit) to
interleaved RGB, as `cv2.cvtColor(yuv, cv2.COLOR_YUV2RGB_I420)` computes it in
process_yuv_video (utils.py:134-156 of the reference).

TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench's CPU legs).

The arithmetic lives in OpenCV (third party, `opencv-python`, unpinned in the reference's
requirements.txt:10; 4.13.0 in this image), not under /root/reference.  Its published algorithm
(modules/imgproc/src/color_yuv.simd.hpp, ITU-R BT.601 limited range in 20-bit fixed point):
    y'  = max(0, Y - 16) * 1220542
    R   = sat8((y' + (1 << 19) + 1673527 * (V - 128)) >> 20)
    G   = sat8((y' + (1 << 19) -  852492 * (V - 128) - 409993 * (U - 128)) >> 20)
    B   = sat8((y' + (1 << 19) + 2116026 * (U - 128)) >> 20)
with U, V of the 2x2 block the pixel belongs to (no chroma interpolation), `>>` arithmetic.
Pinned: tests/golden/make_golden_yuv.py ran cv2 4.13.0 over ALL 2^24 (Y, U, V) triples and a
random frame; tests/test_yuv_cpu.py checks this restatement against the recorded digest/frame."""
import numpy as np

CY, CUB, CUG, CVG, CVR, SHIFT = 1220542, 2116026, -409993, -852492, 1673527, 20


def yuv_to_rgb_values(y, u, v):
    """elementwise on integer arrays (any broadcastable shapes) -> (r, g, b) uint8"""
    y = np.maximum(np.asarray(y, np.int64) - 16, 0) * CY
    uu = np.asarray(u, np.int64) - 128
    vv = np.asarray(v, np.int64) - 128
    half = 1 << (SHIFT - 1)
    r = (y + half + CVR * vv) >> SHIFT
    g = (y + half + CVG * vv + CUG * uu) >> SHIFT
    b = (y + half + CUB * uu) >> SHIFT
    return tuple(np.clip(c, 0, 255).astype(np.uint8) for c in (r, g, b))


def i420_to_rgb(yuv: np.ndarray, height: int, width: int) -> np.ndarray:
    """yuv: uint8 [height*3/2, width] (or flat): Y plane, then U and V planes of (height/2, width/2)
    -> uint8 [height, width, 3] RGB"""
    if height % 2 or width % 2:
        raise ValueError("I420 needs even width and height")
    flat = np.ascontiguousarray(yuv, dtype=np.uint8).reshape(-1)
    if flat.size != height * width * 3 // 2:
        raise ValueError("I420 frame must hold width*height*3/2 bytes")
    hw = height * width
    Y = flat[:hw].reshape(height, width)
    U = flat[hw:hw + hw // 4].reshape(height // 2, width // 2)
    V = flat[hw + hw // 4:].reshape(height // 2, width // 2)
    up = lambda p: np.repeat(np.repeat(p, 2, axis=0), 2, axis=1)
    r, g, b = yuv_to_rgb_values(Y, up(U), up(V))
    return np.stack([r, g, b], axis=-1)


def all_triples_frame():
    """an I420 frame (4096 x 4096) in which every (Y, U, V) triple occurs: block (by, bx) carries
    U = by & 255, V = bx & 255 and its four pixels plus the 8 x 8 block neighbourhood enumerate Y"""
    H = W = 4096
    by, bx = np.meshgrid(np.arange(H // 2), np.arange(W // 2), indexing="ij")
    U = (by & 255).astype(np.uint8)
    V = (bx & 255).astype(np.uint8)
    py, px = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    # blocks repeat every 256 in both directions: (by >> 8, bx >> 8) in 8 x 8, times 4 pixels = 256 Y values
    Y = ((((py >> 9) * 8 + (px >> 9)) << 2) | ((py & 1) << 1) | (px & 1)).astype(np.uint8)
    return np.concatenate([Y.reshape(-1), U.reshape(-1), V.reshape(-1)]), H, W
