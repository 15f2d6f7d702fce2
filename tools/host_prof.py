"""Host-side cost of the eager operator path: cProfile over fit steps in sync-free mode (the GPU never makes the host
wait, so wall time per step = host time per step).  python tools/host_prof.py [C2] [steps]"""
import cProfile
import pstats
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
s = make_config(cfg)
d = s.to("cuda:0")
means = d.means2d.clone().requires_grad_(True)
L = d.cholesky.clone().requires_grad_(True)
colors = d.colors.clone().requires_grad_(True)


def step():
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W)
    loss = torch.nn.functional.mse_loss(out, d.gt)
    return torch.autograd.grad(loss, [means, L, colors])


def render():
    with torch.no_grad():
        xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
        return G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W)


for mode in (False, True):
    G.set_sync_free(mode)
    for name, fn in (("fit", step), ("render", render)):
        for _ in range(20):
            fn()
        best = []
        for _ in range(5):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(steps):
                fn()
            t_host = time.perf_counter() - t0
            torch.cuda.synchronize()
            best.append((time.perf_counter() - t0, t_host))
        t_all, t_host = min(best)
        print(f"{name:6s} sync_free={mode}: host {1e6 * t_host / steps:7.1f} us/step, with drain {1e6 * t_all / steps:7.1f} us/step "
              f"(best of 5; worst {1e6 * max(best)[0] / steps:.1f})")

G.set_sync_free(True)
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)

# ---- where a fit step's host time goes, call by call (sync-free, so no call waits for the GPU) -------------------------
import collections

acc = collections.OrderedDict((k, 0.0) for k in ("project", "rasterize", "mse_loss", "autograd.grad"))
for _ in range(20):
    step()
torch.cuda.synchronize()
for _ in range(steps):
    t0 = time.perf_counter()
    xys, depths, radii, conics, nth = G.project_gaussians_2d(means, L, s.H, s.W, s.tile_bounds)
    t1 = time.perf_counter()
    out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, d.opacity, s.H, s.W)
    t2 = time.perf_counter()
    loss = torch.nn.functional.mse_loss(out, d.gt)
    t3 = time.perf_counter()
    torch.autograd.grad(loss, [means, L, colors])
    t4 = time.perf_counter()
    acc["project"] += t1 - t0; acc["rasterize"] += t2 - t1; acc["mse_loss"] += t3 - t2; acc["autograd.grad"] += t4 - t3
torch.cuda.synchronize()
print("per call, us/step:", {k: round(1e6 * v / steps, 1) for k, v in acc.items()}, "sum", round(1e6 * sum(acc.values()) / steps, 1))

# the autograd engine's nodes on the CPU side (torch.profiler; its own overhead inflates every entry, read the shares)
try:
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU]) as prof:
        for _ in range(50):
            step()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=70))
except Exception as e:  # noqa: BLE001
    print("torch.profiler unavailable:", e)
