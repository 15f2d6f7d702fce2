"""The reference callers' statement sequence on the GPU (the box has no /root/reference; tests/test_reference_callers_cpu.py
imports the real files where they exist).  `_RefShapedCholesky` keeps gaussianimage_cholesky.py:76-124 statement for
statement -- attribute names, `.clone()` of the opacity buffer, positional BLOCK_H / BLOCK_W, `background=` / `return_alpha=`
keywords, three renders of one projection, clamp / view / permute / contiguous, loss_fn('L2') = F.mse_loss, backward, the
psnr read-back, optimizer step, zero_grad(set_to_none=True), scheduler step -- so that what is exercised is the call
pattern of an unchanged caller: the zero-depth tag must survive into rasterize_gaussians_sum (fast path), renders two and
three must find the memoised binning, and the result must equal the reference kernels'."""
import math

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import gaussianimage_b200 as G
from gaussianimage_b200 import rasterize_sum as RS
from gaussianimage_b200.optimizer import Adan
from gaussianimage_b200.synth import make_scene
from gsplat.project_gaussians_2d import project_gaussians_2d                      # the import lines of the reference files
from gsplat.project_gaussians_2d_scale_rot import project_gaussians_2d_scale_rot
from gsplat.rasterize_sum import rasterize_gaussians_sum
import refcuda
from test_gpu_parity import DEV, _np, assert_grad_close

pytestmark = pytest.mark.gpu


class _RefShapedCholesky(nn.Module):
    def __init__(self, s, device):
        super().__init__()
        self.H, self.W, self.BLOCK_W, self.BLOCK_H = s.H, s.W, 16, 16
        self.tile_bounds = ((self.W + self.BLOCK_W - 1) // self.BLOCK_W, (self.H + self.BLOCK_H - 1) // self.BLOCK_H, 1)
        N = s.means2d.shape[0]
        self._xyz = nn.Parameter(torch.atanh(s.means2d.clamp(-0.9999, 0.9999)))
        self._cholesky = nn.Parameter(s.cholesky - torch.tensor([0.5, 0, 0.5]))
        self.register_buffer('_opacity', torch.ones((N, 1)))
        self._features_dc = nn.Parameter(s.colors.clone())
        self.random_colors = torch.rand(N, 3, generator=torch.Generator().manual_seed(9))
        self.register_buffer('background', torch.ones(3))
        self.register_buffer('cholesky_bound', torch.tensor([0.5, 0, 0.5]).view(1, 3))
        self.no_clamp = False
        self.to(device)
        self.optimizer = Adan(self.parameters(), lr=1e-3)
        self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=20000, gamma=0.5)

    @property
    def get_xyz(self):
        return torch.tanh(self._xyz)

    @property
    def get_features(self):
        return self._features_dc

    @property
    def get_cholesky_elements(self):
        return self._cholesky + self.cholesky_bound

    def forward(self, **kwargs):                                                  # gaussianimage_cholesky.py:76-99
        self.xys, depths, self.radii, conics, num_tiles_hit = project_gaussians_2d(self.get_xyz, self.get_cholesky_elements, self.H, self.W, self.tile_bounds)
        colors = self.get_features
        opacities = self._opacity.clone()
        out_img = rasterize_gaussians_sum(self.xys, depths, self.radii, conics, num_tiles_hit,
                colors, opacities, self.H, self.W, self.BLOCK_H, self.BLOCK_W, background=self.background, return_alpha=False)
        if not self.no_clamp:
            out_img = torch.clamp(out_img, 0, 1)
        out_img = out_img.view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()
        geom_colors = self.random_colors.to(self.xys.device) * 0.5
        gauss_img = rasterize_gaussians_sum(self.xys, depths, self.radii, conics, num_tiles_hit,
        geom_colors, opacities, self.H, self.W, self.BLOCK_H, self.BLOCK_W, background=self.background, return_alpha=False)
        gauss_img = torch.clamp(gauss_img, 0, 1)
        gauss_img = gauss_img.view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()
        ones_color = torch.ones_like(colors)
        alpha_img = rasterize_gaussians_sum(self.xys, depths, self.radii, conics, num_tiles_hit,
                ones_color, opacities, self.H, self.W, self.BLOCK_H, self.BLOCK_W, background=self.background, return_alpha=False)
        out_alpha = alpha_img.mean(dim=-1, keepdim=True)
        out_alpha = out_alpha.view(-1, self.H, self.W, 1).permute(0, 3, 1, 2).contiguous()
        return {"render": out_img, "gauss_render": gauss_img, "alpha_map": out_alpha, "final_opacities": opacities}

    def train_iter(self, gt_image, iterations):                                   # gaussianimage_cholesky.py:100-124
        render_pkg = self.forward()
        image = render_pkg["render"]
        loss = F.mse_loss(image, gt_image)                                        # loss_fn(..., 'L2'), utils.py:20-25
        loss.backward()
        with torch.no_grad():
            mse_loss = F.mse_loss(image, gt_image)
            psnr = 10 * math.log10(1.0 / mse_loss.item())
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)
        self.scheduler.step()
        return loss, psnr


def _ref_render(m, s, colors):
    """The same forward through the reference's own kernels and staging (tests/refcuda.py)."""
    with torch.no_grad():
        means, L = m.get_xyz.contiguous(), m.get_cholesky_elements.contiguous()
        xys, depths, radii, conics, nth = refcuda.project_cholesky_fwd(means, L, s.H, s.W)
        n, cum, ids, gids, ids_s, gids_s, bins = refcuda.bin_and_sort(xys, depths, radii, nth, s.H, s.W)
        return refcuda.rasterize_sum_fwd(s.H, s.W, gids_s, bins, xys, conics, colors.contiguous(), m._opacity)[0], xys, radii


@pytest.mark.parametrize("mode", ["default", "python_functions"])
def test_reference_shaped_model_forward_and_train_iter(mode):
    from gaussianimage_b200 import _lib
    s = make_scene(128, 192, 3000, seed=1)
    m = _RefShapedCholesky(s, DEV)
    gt = s.gt.permute(2, 0, 1).unsqueeze(0).contiguous().to(DEV)
    prev = _lib._cpp_autograd
    _lib.set_cpp_autograd(mode == "default" and prev)
    try:
        st0 = RS.stats()
        with torch.no_grad():                                                     # train.py:242-247 (eval FPS loop)
            pkg = m()
        st1 = RS.stats()
        assert st1["fast_calls"] - st0["fast_calls"] == 3, "the zero-depth tag did not survive into rasterize_gaussians_sum"
        assert st1["staged_calls"] == st0["staged_calls"]
        assert st1["geometry_hits"] - st0["geometry_hits"] == 2, "renders two and three must re-use the binning"
        assert m.xys.shape == (3000, 2) and m.radii.dtype == torch.int32          # attributes train.py:184 reads
        if refcuda.available("jit"):
            want, r_xys, r_radii = _ref_render(m, s, m._features_dc)
            assert torch.equal(m.xys, r_xys) and torch.equal(m.radii, r_radii)
            assert torch.equal(pkg["render"], torch.clamp(want, 0, 1).view(-1, s.H, s.W, 3).permute(0, 3, 1, 2))
            want_g = _ref_render(m, s, m.random_colors.to(DEV) * 0.5)[0]
            assert torch.equal(pkg["gauss_render"], torch.clamp(want_g, 0, 1).view(-1, s.H, s.W, 3).permute(0, 3, 1, 2))
            want_a = _ref_render(m, s, torch.ones_like(m._features_dc))[0]
            assert torch.equal(pkg["alpha_map"], want_a.mean(dim=-1, keepdim=True).view(-1, s.H, s.W, 1).permute(0, 3, 1, 2))
        # training iterations in the reference's order; the loss must fall and every parameter must move
        before = [p.detach().clone() for p in m.parameters()]
        losses = []
        for it in range(1, 31):
            loss, psnr = m.train_iter(gt, it)
            losses.append(loss.item())
        assert losses[-1] < losses[0] and math.isfinite(psnr)
        assert all(not torch.equal(a, b.detach()) for a, b in zip(before, m.parameters()))
        G.check_overflow(DEV)
    finally:
        _lib.set_cpp_autograd(prev)


def test_reference_shaped_scale_rot_forward():
    """gaussianimage_rs.py:74-82: project_gaussians_2d_scale_rot(self.get_xyz, self.get_scaling, self.get_rotation, H, W, tile_bounds)
    then ONE rasterize_gaussians_sum with background / return_alpha keywords."""
    s = make_scene(128, 192, 3000, seed=2)
    d = s.to(DEV)
    st0 = RS.stats()
    xys, depths, radii, conics, num_tiles_hit = project_gaussians_2d_scale_rot(d.means2d, d.scales, d.rotation, s.H, s.W, s.tile_bounds)
    out_img = rasterize_gaussians_sum(xys, depths, radii, conics, num_tiles_hit, d.colors, d.opacity, s.H, s.W, 16, 16,
                                      background=torch.ones(3, device=DEV), return_alpha=False)
    assert RS.stats()["fast_calls"] - st0["fast_calls"] == 1
    if refcuda.available("jit"):
        r = refcuda.project_scale_rot_fwd(d.means2d, d.scales, d.rotation, s.H, s.W)
        n, cum, ids, gids, ids_s, gids_s, bins = refcuda.bin_and_sort(r[0], r[1], r[2], r[4], s.H, s.W)
        want = refcuda.rasterize_sum_fwd(s.H, s.W, gids_s, bins, r[0], r[3], d.colors, d.opacity)[0]
        assert torch.equal(out_img, want)
