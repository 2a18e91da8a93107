"""The decode-side restatement (oracle/quant.py) against what the reference's own quantizer classes
produced (tests/golden/quant.npz, tests/golden/make_golden_quant.py)."""
import os

import numpy as np

from oracle import quant as Q

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "quant.npz"))


def test_fp16_means():
    assert np.array_equal(Q.dequantize_xyz(G["xyz_half_bits"]), G["xyz_q"])
    assert np.array_equal(G["xyz"].astype(np.float16).view(np.uint16), G["xyz_half_bits"])     # the encode is a plain cast


def test_cholesky_codes_round_trip():
    assert np.array_equal(Q.dequantize_cholesky(G["chol_code"], G["chol_scale"], G["chol_beta"]), G["chol_deq"])
    code, scale, beta = Q.compress_cholesky(G["chol"])
    assert np.array_equal(scale, G["chol_scale"]) and np.array_equal(beta, G["chol_beta"])
    assert np.array_equal(code, G["chol_code"])
    assert code.min() == 0 and code.max() == 63


def test_residual_vq_decode():
    assert np.array_equal(Q.decode_colours(G["color_index"], G["codebooks"]), G["colours"])
