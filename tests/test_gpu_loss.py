"""GPU: the loss epilogue (rasterize_gaussians_sum_mse) against the sequence the reference's training step
runs -- rasterize_gaussians_sum, clamp, layout change, F.mse_loss (gaussianimage_cholesky.py:80-85,104; utils.py:20-25) --
and against the CPU oracle's fit step.  Loss within 1e-5 relative, gradients within 1e-4 (north_star tolerances)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_scene
from oracle import gi2d_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _np(t):
    return t.detach().cpu().numpy()


def grad_close(got, want, what):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    scale = np.abs(want).max() + 1e-30
    err = np.abs(got - want)
    tol = 1e-4 * np.abs(want) + 1e-5 * scale
    assert (err <= tol).all(), f"{what}: {(err > tol).sum()} beyond 1e-4, max err {err.max():.3e} at scale {scale:.3e}"


def _leaves(d, gain):
    return [d.means2d.clone().requires_grad_(True), d.cholesky.clone().requires_grad_(True),
            (d.colors * gain).clone().requires_grad_(True), d.opacity.clone().requires_grad_(True)]


def _reference_sequence(leaves, s, gt_chw, clamp, upstream):
    means, L, colors, opac = leaves
    pr = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(*pr, colors, opac, s.H, s.W)
    if clamp:
        out = torch.clamp(out, 0, 1)                                              # gaussianimage_cholesky.py:84
    image = out.view(-1, s.H, s.W, 3).permute(0, 3, 1, 2).contiguous()            # :85
    loss = F.mse_loss(image, gt_chw)                                              # utils.py:21
    return loss, torch.autograd.grad(loss * upstream, leaves), image


@pytest.mark.parametrize("H,W,N,dense", [(70, 100, 257, 0.0), (128, 192, 3000, 0.0), (128, 192, 6000, 0.5)])
@pytest.mark.parametrize("clamp", [True, False])
@pytest.mark.parametrize("layout", ["chw", "hwc"])
def test_fused_loss_equals_render_clamp_permute_mse(H, W, N, dense, clamp, layout):
    s = make_scene(H, W, N, seed=1, dense=dense)
    d = s.to(DEV)
    gain = 3.0                       # many pixels above 1: the clamp and its gradient mask are really exercised
    gt_chw = d.gt.permute(2, 0, 1).unsqueeze(0).contiguous()
    a, b = _leaves(d, gain), _leaves(d, gain)
    w_loss, w_grads, w_image = _reference_sequence(a, s, gt_chw, clamp, 1.7)
    pr = G.project_gaussians_2d(b[0], b[1], s.H, s.W, s.tile_bounds)
    loss, image = G.rasterize_gaussians_sum_mse(*pr, b[2], b[3], s.H, s.W, gt_chw if layout == "chw" else d.gt, clamp=clamp,
                                                return_image=True)
    grads = torch.autograd.grad(loss * 1.7, b)
    assert loss.shape == () and abs(loss.item() - w_loss.item()) <= 1e-5 * abs(w_loss.item())
    assert torch.equal(image.permute(2, 0, 1).unsqueeze(0), w_image)
    if clamp:
        assert (w_image == 1).float().mean() > 0.01
    for g_, w_, nm in zip(grads, w_grads, ("means2d", "L", "colors", "opacity")):
        assert g_.shape == w_.shape
        grad_close(_np(g_), _np(w_), f"d loss / d {nm}")


def test_fused_loss_vs_cpu_oracle_fit_step():
    s = make_scene(128, 192, 3000, seed=1)
    d = s.to(DEV)
    leaves = _leaves(d, 1.0)
    pr = G.project_gaussians_2d(leaves[0], leaves[1], s.H, s.W, s.tile_bounds)
    loss = G.rasterize_gaussians_sum_mse(*pr, leaves[2], leaves[3], s.H, s.W, d.gt, clamp=False)
    grads = torch.autograd.grad(loss, leaves)
    w_loss, w = O.fit_step_cholesky(_np(s.means2d), _np(s.cholesky), _np(s.colors), _np(s.opacity), _np(s.gt), s.H, s.W)
    assert abs(loss.item() - w_loss) <= 1e-5 * abs(w_loss)
    for g_, key in zip(grads, ("v_means2d", "v_L", "v_colors", "v_opacity")):
        grad_close(_np(g_), w[key], key)


def test_loss_is_deterministic_and_graph_capturable():
    s = make_scene(128, 192, 3000, seed=1)
    d = s.to(DEV)
    gt_chw = d.gt.permute(2, 0, 1).unsqueeze(0).contiguous()
    leaves = _leaves(d, 1.0)

    def step():
        pr = G.project_gaussians_2d(leaves[0], leaves[1], s.H, s.W, s.tile_bounds)
        loss = G.rasterize_gaussians_sum_mse(*pr, leaves[2], leaves[3], s.H, s.W, gt_chw)
        return loss, torch.autograd.grad(loss, leaves)
    l0, g0 = step()
    l0 = l0.item()                # keep no autograd graph alive across the capture (its AccumulateGrad nodes remember their stream)
    for _ in range(3):
        assert step()[0].item() == l0                 # fixed-order sum of per-CTA partials
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        lg, gg = step()
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    assert lg.item() == l0
    for a, b in zip(gg, g0):
        grad_close(_np(a), _np(b), "graph replay")
    G.check_overflow(DEV)


def test_no_intersections_and_argument_errors():
    H, W = 64, 96
    tb = (6, 4, 1)
    means = torch.full((5, 2), 30.0, device=DEV)               # far off-screen: no tile is hit
    L = torch.tensor([[1.0, 0.0, 1.0]] * 5, device=DEV)
    colors = torch.rand(5, 3, device=DEV).requires_grad_(True)
    opac = torch.ones(5, 1, device=DEV)
    gt = torch.rand(H, W, 3, device=DEV)
    pr = G.project_gaussians_2d(means, L, H, W, tb)
    loss = G.rasterize_gaussians_sum_mse(*pr, colors, opac, H, W, gt)
    want = F.mse_loss(torch.ones(H, W, 3, device=DEV), gt)      # rasterize_sum.py:121-129: ones * background
    assert abs(loss.item() - want.item()) <= 1e-5 * want.item()
    (gc,) = torch.autograd.grad(loss, [colors])
    assert (gc == 0).all()
    with pytest.raises(ValueError):
        G.rasterize_gaussians_sum_mse(*pr, colors, opac, H, W, gt[:-1])
    with pytest.raises(RuntimeError):
        G.rasterize_gaussians_sum_mse(*pr, colors, opac, H, W, gt.cpu())
    with pytest.raises(RuntimeError):
        G.rasterize_gaussians_sum_mse(pr[0], torch.rand(5, device=DEV), *pr[2:], colors, opac, H, W, gt)
