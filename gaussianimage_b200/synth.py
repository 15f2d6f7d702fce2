"""Synthetic Gaussians and target images of the shapes BASELINE.json names (no datasets, no checkpoints).

Parameters are drawn exactly like the reference models initialise theirs (gaussianimage_cholesky.py:30-34,42,74;
gaussianimage_rs.py:26-30,56,60), on the CPU generator so the same seed gives the same tensors on every box.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

CONFIGS = {
    # name: (H, W, N, parametrisation)  -- BASELINE.json configs 1..5 (C1..C5)
    "C1": (512, 768, 10_000, "cholesky"),
    "C2": (512, 768, 70_000, "cholesky"),
    "C3": (678, 1020, 100_000, "rs"),
    "C4": (512, 768, 70_000, "cholesky"),     # x24 images, image-parallel
    "C5": (4320, 7680, 1_000_000, "cholesky"),
}


@dataclass
class Scene:
    H: int
    W: int
    means2d: torch.Tensor        # (N,2) in (-1,1)  = tanh(_xyz)
    cholesky: torch.Tensor       # (N,3) = _cholesky + [0.5, 0, 0.5]            (Cholesky models)
    scales: torch.Tensor         # (N,2) = |_scaling + 0.5|                    (RS model)
    rotation: torch.Tensor       # (N,1) = sigmoid(_rotation) * 2*pi           (RS model)
    colors: torch.Tensor         # (N,3) = _features_dc
    opacity: torch.Tensor        # (N,1) ones
    gt: torch.Tensor             # (H,W,3) smooth target image in [0,1]

    @property
    def tile_bounds(self):
        return ((self.W + 15) // 16, (self.H + 15) // 16, 1)

    def to(self, device):
        return Scene(self.H, self.W, *[t.to(device) for t in (self.means2d, self.cholesky, self.scales, self.rotation,
                                                                self.colors, self.opacity, self.gt)])


def make_scene(H: int, W: int, N: int, seed: int = 1, dense: float = 0.0) -> Scene:
    """`dense` > 0 concentrates that fraction of the Gaussians in a 40-pixel patch so that some tiles exceed the
    reference's 256-intersection cap (quirk Q1)."""
    g = torch.Generator().manual_seed(seed)
    xyz = torch.atanh(2 * (torch.rand(N, 2, generator=g) - 0.5))
    means = torch.tanh(xyz)
    if dense > 0:
        k = int(N * dense)
        means[:k] = (torch.rand(k, 2, generator=g) - 0.5) * torch.tensor([80.0 / W, 80.0 / H]) + 0.1
    chol = torch.rand(N, 3, generator=g) + torch.tensor([0.5, 0.0, 0.5])
    colors = torch.rand(N, 3, generator=g)
    scales = torch.abs(torch.rand(N, 2, generator=g) + 0.5)
    rotation = torch.sigmoid(torch.rand(N, 1, generator=g)) * 2 * math.pi
    opacity = torch.ones(N, 1)
    g2 = torch.Generator().manual_seed(seed + 1)
    low = torch.rand(1, 3, (H + 15) // 16 + 1, (W + 15) // 16 + 1, generator=g2)
    gt = torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)[0].permute(1, 2, 0)
    gt = torch.round(gt * 255).clamp(0, 255).to(torch.uint8).float() / 255     # 8-bit levels, like a decoded Kodak/DIV2K image
    return Scene(H, W, means.contiguous(), chol.contiguous(), scales.contiguous(), rotation.contiguous(),
                 colors.contiguous(), opacity, gt.contiguous())


def make_config(name: str, seed: int = 1) -> Scene:
    H, W, N, _ = CONFIGS[name]
    return make_scene(H, W, N, seed)
