"""CPU: the decode-prologue oracle against the reference's own quantizer decompress methods (tests/golden/decode_reference.npz,
made by tests/golden/make_golden_decode.py).  Cholesky and colours bit-exact; tanh within 1 ulp of torch's (libm vs SLEEF)."""
import numpy as np

from oracle.decode_oracle import decode_prologue

G = np.load("tests/golden/decode_reference.npz")


def test_decode_prologue_oracle_matches_reference_quantizers():
    means, chol, colors = decode_prologue(G["xyz"], G["codes"], G["scale"], G["beta"], G["bound"], G["index"], G["codebooks"])
    assert np.array_equal(chol.view(np.int32), G["cholesky"].view(np.int32))
    assert np.array_equal(colors.view(np.int32), G["colors"].view(np.int32))
    np.testing.assert_allclose(means, G["means"], rtol=2e-7, atol=0)
    assert means.dtype == np.float32 and np.abs(means).max() < 1


def test_uint8_codes_decode_like_float_codes():
    a = decode_prologue(G["xyz"], G["codes"], G["scale"], G["beta"], G["bound"], G["index"], G["codebooks"])
    b = decode_prologue(G["xyz"], G["codes"].astype(np.uint8), G["scale"], G["beta"], G["bound"], G["index"].astype(np.int32), G["codebooks"])
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
