"""The I420 -> RGB restatement (oracle/yuv.py) against what OpenCV -- the library the reference's
loader calls, utils.py:150 -- produced in the build container (tests/golden/yuv.npz)."""
import hashlib
import os

import numpy as np
import pytest

from oracle import yuv as Y

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "yuv.npz"))


def test_all_2_pow_24_triples_match_opencv():
    frame, H, W = Y.all_triples_frame()
    hw = H * W
    up = lambda p: np.repeat(np.repeat(p.reshape(H // 2, W // 2), 2, 0), 2, 1).astype(np.uint32)
    t = (frame[:hw].reshape(H, W).astype(np.uint32) << 16) | (up(frame[hw:hw + hw // 4]) << 8) | up(frame[hw + hw // 4:])
    assert np.unique(t).size == 1 << 24                      # the frame really holds every (Y, U, V)
    rgb = Y.i420_to_rgb(frame, H, W)
    assert hashlib.sha256(np.ascontiguousarray(rgb).tobytes()).hexdigest() == str(G["all_triples_sha256"])


def test_small_frame_matches_opencv():
    assert np.array_equal(Y.i420_to_rgb(G["small_yuv"], 48, 64), G["small_rgb"])


def test_live_opencv_if_present():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(11)
    h, w = 270, 480
    f = rng.integers(0, 256, size=h * w * 3 // 2, dtype=np.uint8)
    assert np.array_equal(Y.i420_to_rgb(f, h, w), cv2.cvtColor(f.reshape(h * 3 // 2, w), cv2.COLOR_YUV2RGB_I420))


def test_shape_errors():
    with pytest.raises(ValueError):
        Y.i420_to_rgb(np.zeros(10, np.uint8), 3, 4)
    with pytest.raises(ValueError):
        Y.i420_to_rgb(np.zeros(12, np.uint8), 4, 4)
