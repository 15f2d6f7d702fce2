"""GPU: the fused decode-prologue + projection kernel (gi2d_decode_project_cholesky_fwd) against
 (a) the reference's sequence of torch ops on the GPU followed by the (parity-pinned) projection operator -- bit-exact,
 (b) the CPU oracle on the golden representation of the reference's quantizers."""
import numpy as np
import pytest
import torch

import gaussianimage_b200 as G
from oracle import gi2d_oracle as O
from oracle.decode_oracle import decode_prologue

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = np.load("tests/golden/decode_reference.npz")


def _torch_sequence(xyz, codes, scale, beta, bound, index, cb, H, W, tb):
    means = torch.tanh(xyz.float())                                     # gaussianimage_cholesky.py:163
    chol = codes.float() * scale + beta + bound                         # quantize.py:87 + :165
    colors = 0                                                          # quantize.py:147-149
    for k in range(index.shape[1]):
        colors = colors + cb[k][index[:, k].long()]
    return (*G.project_gaussians_2d(means, chol.contiguous(), H, W, tb), colors)


@pytest.mark.parametrize("code_dtype", [torch.float32, torch.uint8])
@pytest.mark.parametrize("idx_dtype", [torch.int64, torch.int32])
def test_fused_decode_equals_torch_ops_then_projection_bit_for_bit(code_dtype, idx_dtype):
    H, W = 300, 420
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    t = {k: torch.from_numpy(GOLD[k]).to(DEV) for k in ("xyz", "codes", "scale", "beta", "bound", "index", "codebooks")}
    codes, index = t["codes"].to(code_dtype), t["index"].to(idx_dtype)
    got = G.decode_project_gaussians_2d(t["xyz"], codes, t["scale"], t["beta"], t["bound"], index, t["codebooks"], H, W, tb)
    want = _torch_sequence(t["xyz"], codes, t["scale"], t["beta"], t["bound"], index, t["codebooks"], H, W, tb)
    for g, w, nm in zip(got, want, ("xys", "depths", "radii", "conics", "num_tiles_hit", "colors")):
        assert g.shape == w.shape and g.dtype == w.dtype, nm
        assert torch.equal(g, w), f"{nm} differs from torch ops + projection"
    # and a render of the fused outputs takes the fused (zero-depth) path and equals the render of the sequence's outputs
    op = torch.ones(got[0].shape[0], 1, device=DEV)
    img_a = G.rasterize_gaussians_sum(*got[:5], got[5], op, H, W)
    img_b = G.rasterize_gaussians_sum(*want[:5], want[5], op, H, W)
    assert torch.equal(img_a, img_b)


def test_fused_decode_vs_cpu_oracle():
    H, W = 300, 420
    tb = ((W + 15) // 16, (H + 15) // 16, 1)
    t = {k: torch.from_numpy(GOLD[k]).to(DEV) for k in ("xyz", "codes", "scale", "beta", "bound", "index", "codebooks")}
    xys, depths, radii, conics, nth, colors = G.decode_project_gaussians_2d(t["xyz"], t["codes"], t["scale"], t["beta"], t["bound"],
                                                                         t["index"], t["codebooks"], H, W, tb)
    means, chol, cols = decode_prologue(GOLD["xyz"], GOLD["codes"], GOLD["scale"], GOLD["beta"], GOLD["bound"], GOLD["index"],
                                        GOLD["codebooks"])
    assert np.array_equal(colors.cpu().numpy().view(np.int32), cols.view(np.int32))
    w_xys, w_depths, w_radii, w_conics, w_nth = O.project_gaussians_2d(means, chol, H, W)
    # positions go through tanhf (libdevice on the GPU, libm in the oracle: <= 1 ulp apart); covariances do not
    np.testing.assert_allclose(xys.cpu().numpy(), w_xys, rtol=1e-6, atol=1e-4)
    assert np.array_equal(radii.cpu().numpy(), w_radii)
    assert np.array_equal(conics.cpu().numpy().view(np.int32), w_conics.view(np.int32))
    assert (nth.cpu().numpy() != w_nth).mean() < 5e-3
    assert (depths == 0).all()


def test_ragged_sizes_and_errors():
    H, W = 64, 96
    tb = (6, 4, 1)
    g = torch.Generator().manual_seed(1)
    for N in (1, 255, 257):
        xyz = (torch.rand(N, 2, generator=g) * 2 - 1).half().to(DEV)
        codes = torch.randint(0, 64, (N, 3), generator=g).to(torch.uint8).to(DEV)
        idx = torch.randint(0, 8, (N, 2), generator=g).to(DEV)
        cb = torch.rand(2, 8, 3, generator=g).to(DEV)
        sc, be, bo = torch.full((3,), 1 / 63, device=DEV), torch.zeros(3, device=DEV), torch.tensor([[0.5, 0.0, 0.5]], device=DEV)
        got = G.decode_project_gaussians_2d(xyz, codes, sc, be, bo, idx, cb, H, W, tb)
        want = _torch_sequence(xyz, codes, sc, be, bo.view(1, 3), idx, cb, H, W, tb)
        for a, b in zip(got, want):
            assert torch.equal(a, b)
    with pytest.raises(RuntimeError):
        G.decode_project_gaussians_2d(xyz.float(), codes, sc, be, bo, idx, cb, H, W, tb)
    with pytest.raises(RuntimeError):
        G.decode_project_gaussians_2d(xyz.cpu(), codes, sc, be, bo, idx, cb, H, W, tb)
    with pytest.raises(RuntimeError):
        G.decode_project_gaussians_2d(xyz, codes, sc, be, bo, idx[:, :1], cb, H, W, tb)
