"""Per-kernel timeline of the CUDA-graph replay of one fit step (CUPTI via torch.profiler): start offset, duration, gap.
python tools/graph_timeline.py [C2]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
s = make_config(cfg)
d = s.to("cuda:0")
means = d.means2d.clone().requires_grad_(True)
L = d.cholesky.clone().requires_grad_(True)
colors = d.colors.clone().requires_grad_(True)
bg = torch.ones(3, device="cuda:0")


def step():
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W, background=bg)
    loss = torch.nn.functional.mse_loss(out, d.gt)
    return torch.autograd.grad(loss, [means, L, colors])


side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        step()
torch.cuda.synchronize()
from gaussianimage_b200 import graphs as GR
g = GR.new_graph()
with torch.cuda.graph(g):
    keep = step()
print("memset nodes replaced by fill kernels:", GR.fold_memsets(g))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for _ in range(5):
    g.replay()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        flush.zero_()
        g.replay()
    torch.cuda.synchronize()
evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
# last replay = events after the last flush fill
idx = max(i for i, e in enumerate(evs) if "FillFunctor<unsigned char>" in e.name or ("Memset" in e.name and e.time_range.elapsed_us() > 20))
run = evs[idx + 1:]
t0 = run[0].time_range.start
prev_end = t0
for e in run:
    st, en = e.time_range.start, e.time_range.end
    print(f"+{st - t0:8.1f} us  dur {en - st:7.1f}  gap {st - prev_end:6.1f}  {e.name[:90]}")
    prev_end = en
print(f"total {prev_end - t0:.1f} us")
