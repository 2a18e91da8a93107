"""CPU restatement of the codec's decode arithmetic (wire format -> the Gaussians the two operators
see), following the reference's quantize.py.  TEST INFRASTRUCTURE ONLY.

  dequantize_xyz       FakeQuantizationHalf.forward, quantize.py:15-20: the fp16 value as fp32
  dequantize_cholesky  UniformQuantizer.decompress, quantize.py:86-88: code * scale + beta (two fp32
                       roundings), per channel
  decode_colours       VectorQuantizer.decompress, quantize.py:146-150: 0 + e0[i0] + e1[i1]
  compress_cholesky    UniformQuantizer._init_data + compress, quantize.py:41-47, :82-84 (to make
                       wire data for tests; the decode never needs it)
Pinned by tests/test_quant_cpu.py against tests/golden/quant.npz, which the reference's own classes
produced (tests/golden/make_golden_quant.py)."""
import numpy as np

_f = np.float32


def dequantize_xyz(half_bits):
    return np.asarray(half_bits, np.uint16).view(np.float16).astype(_f)


def dequantize_cholesky(code, scale, beta):
    return (np.asarray(code).astype(_f) * np.asarray(scale, _f)).astype(_f) + np.asarray(beta, _f)


def decode_colours(index, codebooks):
    cb = np.asarray(codebooks, _f)
    ix = np.asarray(index).astype(np.int64)
    return ((_f(0) + cb[0][ix[:, 0]]).astype(_f) + cb[1][ix[:, 1]]).astype(_f)


def compress_cholesky(chol, bits=6):
    chol = np.asarray(chol, _f)
    qmax = 2 ** bits - 1
    beta = chol.min(axis=0)
    scale = ((chol.max(axis=0) - beta) / _f(qmax)).astype(_f)
    code = np.clip(((chol - beta) / scale).astype(_f), 0, qmax)
    return np.round(code).astype(np.uint8), scale, beta        # np.round = round-half-to-even, as torch.round
