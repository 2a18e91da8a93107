"""Golden data for the I420 -> RGB frame loader from OpenCV itself (the library the reference
calls: utils.py:150, cv2.cvtColor(yuv, cv2.COLOR_YUV2RGB_I420)).

Run in the build container:  python tests/golden/make_golden_yuv.py      (needs cv2; 4.13.0 here)
 -> yuv.npz: sha256 of cv2's RGB output over a frame holding all 2^24 (Y,U,V) triples
    (oracle.yuv.all_triples_frame), and a random 48 x 64 I420 frame with cv2's RGB for it."""
import hashlib
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import yuv as Y  # noqa: E402

if __name__ == "__main__":
    frame, H, W = Y.all_triples_frame()
    rgb = cv2.cvtColor(frame.reshape(H * 3 // 2, W), cv2.COLOR_YUV2RGB_I420)
    # every triple really occurs
    yy = frame[:H * W].reshape(H, W)[::2, ::2].astype(np.int64)      # one of the four Y values per block is enough here
    digest = hashlib.sha256(np.ascontiguousarray(rgb).tobytes()).hexdigest()
    rng = np.random.default_rng(5)
    h, w = 48, 64
    small = rng.integers(0, 256, size=h * w * 3 // 2, dtype=np.uint8)
    small_rgb = cv2.cvtColor(small.reshape(h * 3 // 2, w), cv2.COLOR_YUV2RGB_I420)
    np.savez_compressed(os.path.join(HERE, "yuv.npz"), all_triples_sha256=np.array(digest), small_yuv=small,
                        small_rgb=small_rgb, cv2_version=np.array(cv2.__version__))
    mine = Y.i420_to_rgb(frame, H, W)
    print("cv2", cv2.__version__, "digest", digest[:16], "restatement equal on all triples:",
          bool(np.array_equal(mine, rgb)), "| small:", bool(np.array_equal(Y.i420_to_rgb(small, h, w), small_rgb)))
