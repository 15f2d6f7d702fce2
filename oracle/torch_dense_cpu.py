"""TEST / BENCH INFRASTRUCTURE ONLY -- never imported by the product (gaussianimage_b200/).

BASELINE.json configs[0]: "GaussianImage_Cholesky fwd+bwd ... via gsplat's pure-PyTorch _torch_impl projection + dense
per-pixel accumulated sum on CPU (ref path)" -- the CPU figure north_star wants beside every GPU number.  The reference
has no such function as a whole (`_torch_impl.py` holds no 2D projection and no accumulated-sum rasterizer),
so this file composes it from what the reference does have, in torch ops on the CPU:

* Cholesky -> covariance and pixel centre as the kernel states them (gsplat/gsplat/cuda/csrc/foward2d.cu:36-47):
  centre = 0.5*W*mx + 0.5*W, cov = (l11^2, l11*l21, l21^2 + l22^2);
* conic and radius by the reference's own `_torch_impl.compute_cov2d_bounds` (gsplat/gsplat/_torch_impl.py:197-212,
  restated below because /root/reference does not exist on the GPU box: determinant CLAMPED to 1e-6, not rejected);
* the accumulated sum of the CUDA rasterizer (gsplat/gsplat/cuda/csrc/forward.cu:591-612) evaluated DENSELY: every
  Gaussian at every pixel, sigma = 0.5*(A dx^2 + C dy^2) + B dx dy, alpha = min(1, opac*exp(-sigma)), pairs with
  sigma < 0 or alpha < 1/255 masked out; no tiles, no 3-sigma bounding box, no 256-per-tile cap (quirk Q1);
* MSE loss and torch autograd for the backward.

Dense and tile-based sums differ where a pair outside the 3-sigma tile box still passes alpha >= 1/255 (possible when
opacity > 255*exp(-4.5) = 0.35) and in tiles with more than 256 intersections; below both limits they agree, which is how
tests/test_oracle_cpu.py ties this path to the oracle.  Memory is bounded by chunking over Gaussians and by a two-pass
backward (image first, then d loss / d image pushed through each chunk again).
"""
from __future__ import annotations

import time

import torch


def compute_cov2d_bounds(cov2d: torch.Tensor, eps: float = 1e-6):
    """gsplat/gsplat/_torch_impl.py:197-212 on upper-triangle (N,3) input: (conic (N,3), radius (N,), valid (N,))."""
    a, b, c = cov2d[..., 0], cov2d[..., 1], cov2d[..., 2]
    det = torch.clamp(a * c - b ** 2, min=eps)
    conic = torch.stack([c / det, -b / det, a / det], dim=-1)
    m = (a + c) / 2
    root = torch.sqrt(torch.clamp(m ** 2 - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(torch.max(m + root, m - root)))
    return conic, radius, (a * c - b ** 2) > eps


def project_cholesky(means2d: torch.Tensor, L: torch.Tensor, H: int, W: int):
    """foward2d.cu:36-47 in torch ops -> (xys (N,2), conics (N,3), radii (N,))."""
    xys = torch.stack([0.5 * W * means2d[:, 0] + 0.5 * W, 0.5 * H * means2d[:, 1] + 0.5 * H], dim=-1)
    cov = torch.stack([L[:, 0] * L[:, 0], L[:, 0] * L[:, 1], L[:, 1] * L[:, 1] + L[:, 2] * L[:, 2]], dim=-1)
    conics, radii, _ = compute_cov2d_bounds(cov)
    return xys, conics, radii


def _chunk_sum(xys, conics, colors, opacity, px, py):
    """Accumulated sum of one chunk of Gaussians over the whole pixel grid (forward.cu:591-612): (H,W,3)."""
    dx = xys[:, 0].view(1, 1, -1) - px                      # (H,W,G)
    dy = xys[:, 1].view(1, 1, -1) - py
    sigma = 0.5 * (conics[:, 0] * dx * dx + conics[:, 2] * dy * dy) + conics[:, 1] * dx * dy
    alpha = torch.clamp(opacity.view(1, 1, -1) * torch.exp(-sigma), max=1.0)
    alpha = torch.where((sigma >= 0) & (alpha >= 1.0 / 255.0), alpha, torch.zeros((), dtype=alpha.dtype))
    return alpha @ colors                                   # (H,W,G) @ (G,3)


def render_dense(means2d, L, colors, opacity, H, W, chunk=64):
    """Forward only: out_img (H,W,3) fp32."""
    py, px = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    px, py = px.unsqueeze(-1), py.unsqueeze(-1)
    out = torch.zeros(H, W, 3)
    with torch.no_grad():
        xys, conics, _ = project_cholesky(means2d, L, H, W)
        for s in range(0, means2d.shape[0], chunk):
            e = s + chunk
            out += _chunk_sum(xys[s:e], conics[s:e], colors[s:e], opacity[s:e].reshape(-1), px, py)
    return out


def fit_step_dense(means2d, L, colors, opacity, gt, H, W, chunk=64):
    """One fwd+bwd fit step: (loss, v_means2d, v_L, v_colors), gradients by torch autograd, chunk by chunk."""
    out = render_dense(means2d, L, colors, opacity, H, W, chunk)
    loss = torch.mean((out - gt) ** 2)
    v_out = 2.0 * (out - gt) / out.numel()                  # d mse / d out
    py, px = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    px, py = px.unsqueeze(-1), py.unsqueeze(-1)
    v_m, v_L, v_c = torch.zeros_like(means2d), torch.zeros_like(L), torch.zeros_like(colors)
    for s in range(0, means2d.shape[0], chunk):
        e = s + chunk
        m, l, c = (t[s:e].detach().clone().requires_grad_(True) for t in (means2d, L, colors))
        xys, conics, _ = project_cholesky(m, l, H, W)
        part = _chunk_sum(xys, conics, c, opacity[s:e].reshape(-1), px, py)
        g = torch.autograd.grad((part * v_out).sum(), [m, l, c])
        v_m[s:e], v_L[s:e], v_c[s:e] = g
    return float(loss), v_m, v_L, v_c


def timed_sample(scene, workload="fit", n_sample=512, chunk=64, threads=None):
    """Time the dense path on the first `n_sample` Gaussians of `scene` over the full image and scale to all N (its cost
    is linear in N: every Gaussian meets every pixel).  Returns (steps_per_s for the full workload, seconds measured)."""
    if threads:
        torch.set_num_threads(threads)
    n = min(n_sample, scene.means2d.shape[0])
    a = [t[:n].contiguous() for t in (scene.means2d, scene.cholesky, scene.colors, scene.opacity)]
    if workload == "fit":
        fn = lambda: fit_step_dense(*a, scene.gt, scene.H, scene.W, chunk)
    else:
        fn = lambda: render_dense(*a, scene.H, scene.W, chunk)
    fn_warm = a[0][:chunk], a[1][:chunk], a[2][:chunk], a[3][:chunk]
    render_dense(*fn_warm, scene.H, scene.W, chunk)         # warm-up: thread pool, allocator
    t0 = time.perf_counter()
    fn()
    dt = time.perf_counter() - t0
    return 1.0 / (dt * scene.means2d.shape[0] / n), dt
