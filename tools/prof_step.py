"""Small driver for ncu: a few eager fit steps (or renders) of one configuration.  python tools/prof_step.py [C2] [steps]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
render_only = len(sys.argv) > 3 and sys.argv[3] == "render"
s = make_config(cfg)
d = s.to("cuda:0")
means = d.means2d.clone().requires_grad_(True)
L = d.cholesky.clone().requires_grad_(True)
colors = d.colors.clone().requires_grad_(True)
for _ in range(steps):
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W)
    if not render_only:
        loss = torch.nn.functional.mse_loss(out, d.gt)
        loss.backward()
torch.cuda.synchronize()
print("ok", float(out.sum()))
