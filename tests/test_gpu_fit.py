"""GPU: the whole training iteration of the reference (gaussianimage_cholesky.py:100-125) as one CUDA graph
(gaussianimage_b200/fit.py: projection, rasterizer with the loss epilogue, backward, capturable Adan) -- it must actually
fit an image, and it must track the same iteration composed from the reference-API operators and the foreach Adan."""
import numpy as np
import pytest
import torch

from gaussianimage_b200.fit import GraphFitter
from gaussianimage_b200.synth import make_scene

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _target(H, W, seed):
    return make_scene(H, W, 1, seed=seed).gt.permute(2, 0, 1).unsqueeze(0).contiguous().to(DEV)


def test_graph_replayed_training_fits_the_image():
    gt = _target(128, 192, 7)
    f = GraphFitter(gt, 3000, lr=0.018, seed=1)
    f.capture()
    f.run(1)
    p0 = f.psnr()
    f.run(600)
    p1 = f.psnr()
    assert np.isfinite(p1) and p1 > p0 + 8.0, (p0, p1)
    assert f.opt.device_step() == f.iteration == 602
    # the learning-rate schedule reaches the device-resident hyper-parameters
    f.lr_step = 700
    f.run(200)
    assert abs(f.opt.param_groups[0]["lr"] - 0.009) < 1e-12
    assert f.psnr() > p1 - 0.5


def test_fused_iteration_tracks_the_reference_api_iteration():
    gt = _target(96, 128, 3)
    a = GraphFitter(gt, 1500, seed=2, fused=True)
    b = GraphFitter(gt, 1500, seed=2, fused=False)
    for it in range(12):
        la, lb = a.step().item(), b.step().item()
        a.iteration += 1
        b.iteration += 1
        assert abs(la - lb) <= 2e-4 * abs(lb), (it, la, lb)       # fp32 atomics order differs between the two backward passes
    for pa, pb in zip(a.params, b.params):
        torch.testing.assert_close(pa, pb, rtol=0, atol=2e-3)


def test_folding_memset_nodes_of_a_captured_step_changes_nothing_but_the_node_kinds():
    """graphs.fold_memsets: the memset node that F.mse_loss's reduction records becomes a fill kernel; loss and gradients of the
    replayed step stay bit-identical to the unedited graph's."""
    import torch.nn.functional as F

    import gaussianimage_b200 as G
    from gaussianimage_b200 import graphs as GR
    from gaussianimage_b200.synth import make_scene

    s = make_scene(512, 768, 20000, seed=5)          # large enough for ATen's mean to take its multi-block path (semaphore memset)
    d = s.to("cuda:0")
    means = d.means2d.clone().requires_grad_(True)
    L = d.cholesky.clone().requires_grad_(True)
    colors = d.colors.clone().requires_grad_(True)

    def step():
        pr = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
        out = G.rasterize_gaussians_sum(*pr, colors, d.opacity, s.H, s.W)
        loss = F.mse_loss(out, d.gt)
        return loss, torch.autograd.grad(loss, [means, L, colors])

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    results = []
    for fold in (False, True):
        g = torch.cuda.CUDAGraph(keep_graph=True)
        with torch.cuda.graph(g):
            loss, grads = step()
        n = GR.fold_memsets(g) if fold else 0
        if fold:
            assert n >= 1, "the captured step holds no memset node any more: the helper has nothing to do"
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        results.append((loss.clone(), [x.clone() for x in grads]))
        del g
    assert torch.equal(results[0][0], results[1][0])
    # the rasterizer's gradients are summed with red.global across tiles (order varies run to run): compare at the gradient tolerance
    for a, b in zip(results[0][1], results[1][1]):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5 * float(a.abs().max()))
