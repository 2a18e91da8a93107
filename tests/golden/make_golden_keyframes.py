"""Golden data for the key-frame decision from the reference's own utils.detect_outliers_mean_diff
(utils.py:214-229) and the normalisation of train_video_Represent.py:336-349.
Run in the build container:  python tests/golden/make_golden_keyframes.py  -> keyframes.npz
(utils.py imports pytorch_msssim, absent here and not used by the function: stubbed)."""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

if __name__ == "__main__":
    m = types.ModuleType("pytorch_msssim")
    m.ms_ssim = m.ssim = None
    sys.modules["pytorch_msssim"] = m
    sys.path.insert(0, "/root/reference")
    import utils as U
    rng = np.random.default_rng(4)
    out = {}
    for case in range(6):
        n = [120, 120, 31, 12, 300, 2][case]
        s = np.abs(rng.normal(0.0, 1e-4, size=n))
        s[0] = 0.0
        for k in rng.choice(np.arange(1, n), size=min(n - 1, [4, 0, 2, 1, 9, 1][case]), replace=False):
            s[k] += rng.uniform(2e-3, 8e-3)                      # scene cuts: the P probe is much worse
        if case == 3:
            s[1:] = s[1]                                         # degenerate: all equal (0/0 in the reference -> nan)
        rest = s[1:]
        with np.errstate(all="ignore"):
            norm = [s[0]] + [(v - rest.min()) / (rest.max() - rest.min()) for v in rest]
        ks = [int(x + 1) for x in U.detect_outliers_mean_diff(norm)]      # 1-based, :348
        ks = np.insert(ks, 0, 1).astype(int)                              # :349-350
        out[f"scores{case}"] = s
        out[f"kframes{case}"] = ks
        print(case, n, ks)
    np.savez_compressed(os.path.join(HERE, "keyframes.npz"), **out)
