"""TEST INFRASTRUCTURE -- CPU restatement (numpy fp32) of the codec decode prologue of decompress_wo_ec
(gaussianimage_cholesky.py:161-166): tanh of the fp16 positions, UniformQuantizer.decompress (quantize.py:86-87) plus the
Cholesky bound, VectorQuantizer.decompress (quantize.py:146-150).  Only tests/, smoke() and bench CPU legs may import this.

Parity pinning status: PINNED -- tests/golden/decode_reference.npz holds outputs of the reference's own
UniformQuantizer.decompress / VectorQuantizer.decompress (quantize.py imported from /root/reference with its absent
third-party imports -- vector_quantize_pytorch, constriction, pytorch_msssim -- stubbed; they are not on this path) and
torch.tanh; tests/test_decode_cpu.py checks this file against them.
"""
import numpy as np

f32 = np.float32


def decode_prologue(xyz_half, codes, scale, beta, cholesky_bound, index, codebooks):
    """Returns (means (N,2), cholesky (N,3), colors (N,3)), fp32, roundings per torch op."""
    means = np.tanh(np.asarray(xyz_half, dtype=np.float16).astype(f32)).astype(f32)          # gaussianimage_cholesky.py:163
    q = np.asarray(codes).astype(f32)
    chol = (q * np.asarray(scale, f32)).astype(f32) + np.asarray(beta, f32)                    # quantize.py:87
    chol = (chol.astype(f32) + np.asarray(cholesky_bound, f32).reshape(1, 3)).astype(f32)     # gaussianimage_cholesky.py:165
    index = np.asarray(index)
    cb = np.asarray(codebooks, f32)
    colors = np.zeros((index.shape[0], 3), f32)                                                # quantize.py:147: recon = 0
    for k in range(index.shape[1]):
        colors = (colors + cb[k][index[:, k]]).astype(f32)                                     # quantize.py:149
    return means, chol, colors
