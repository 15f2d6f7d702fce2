"""Which torch ops launch which kernels in one eager fit step (torch.profiler table)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile
import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config
s = make_config("C2"); d = s.to("cuda:0")
means = d.means2d.clone().requires_grad_(True); L = d.cholesky.clone().requires_grad_(True); colors = d.colors.clone().requires_grad_(True)
bg = torch.ones(3, device="cuda:0")
def step():
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W, background=bg)
    loss = torch.nn.functional.mse_loss(out, d.gt)
    return torch.autograd.grad(loss, [means, L, colors])
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
