"""Per-kernel GPU time of the render (or fit) step via CUPTI (torch.profiler), averaged over reps; L2 flushed between steps.
python tools/kernel_times.py [C5] [render|fit] [reps]"""
import collections
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C5"
mode = sys.argv[2] if len(sys.argv) > 2 else "render"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
s = make_config(cfg)
d = s.to("cuda:0")
means = d.means2d.clone().requires_grad_(mode == "fit")
L = d.cholesky.clone().requires_grad_(mode == "fit")
colors = d.colors.clone().requires_grad_(mode == "fit")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
G.set_sync_free(True)


def step():
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W)
    if mode == "fit":
        torch.autograd.grad(torch.nn.functional.mse_loss(out, d.gt), [means, L, colors])


with torch.set_grad_enabled(mode == "fit"):
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            flush.zero_()
            step()
        torch.cuda.synchronize()
agg = collections.defaultdict(list)
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        agg[e.name[:90]].append(e.device_time)
tot = 0.0
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    if "FillFunctor<unsigned char>" in k:
        continue
    print(f"{sum(v) / reps:9.1f} us/step  x{len(v) / reps:4.1f}  {k}")
    tot += sum(v) / reps
print(f"{tot:9.1f} us/step total (kernels only, flush excluded)")
