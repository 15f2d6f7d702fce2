"""Driver for one `ncu --set full --profile-from-start off` capture of EVERY kernel of the path: warm-up iterations run
unprofiled, then exactly one iteration between cudaProfilerStart/Stop.
  python tools/prof_all.py c2    (or c1 / c3) fit step (project, bin_small, raster fwd/bwd, project bwd) + loss-epilogue step + Adan step
  python tools/prof_all.py c5    8K decode from the compressed representation (decode_project, scatter binning, raster fwd)
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200.optimizer import Adan
from gaussianimage_b200.synth import make_config

mode = sys.argv[1] if len(sys.argv) > 1 else "c2"
dev = "cuda:0"
G.set_sync_free(True)
if mode in ("c1", "c2", "c3"):
    s = make_config(mode.upper())
    d = s.to(dev)
    ps = [torch.nn.Parameter(x.clone()) for x in (d.means2d, d.cholesky, d.colors)]
    opt = Adan(ps, lr=1e-3)

    def it():
        for fused in (False, True):
            for p in ps:
                p.grad = None
            pr = G.project_gaussians_2d(ps[0], ps[1], s.H, s.W, s.tile_bounds)
            if fused:
                loss = G.rasterize_gaussians_sum_mse(*pr, ps[2], d.opacity, s.H, s.W, d.gt, clamp=True)
            else:
                out = G.rasterize_gaussians_sum(*pr, ps[2], d.opacity, s.H, s.W)
                G.rasterize_gaussians_sum(*pr, torch.ones_like(ps[2]), d.opacity, s.H, s.W)      # memoised: repack only
                loss = torch.nn.functional.mse_loss(out, d.gt)
            loss.backward()
        opt.step()
else:
    s = make_config("C5")
    d = s.to(dev)
    N = d.means2d.shape[0]
    g = torch.Generator().manual_seed(3)
    xyz = torch.atanh(d.means2d.clamp(-0.9995, 0.9995)).half()
    scale, beta = torch.full((3,), 1.0 / 63, device=dev), torch.zeros(3, device=dev)
    bound = torch.tensor([0.5, 0.0, 0.5], device=dev)
    codes = torch.round((d.cholesky - bound) / scale).clamp(0, 63).to(torch.uint8)
    cb = (torch.rand(2, 8, 3, generator=g) * 0.5).to(dev)
    idx = torch.randint(0, 8, (N, 2), generator=g).to(dev)

    def it():
        with torch.no_grad():
            pr = G.decode_project_gaussians_2d(xyz, codes, scale, beta, bound, idx, cb, s.H, s.W, s.tile_bounds)
            G.rasterize_gaussians_sum(*pr[:5], pr[5], d.opacity, s.H, s.W)

for _ in range(3):
    it()
torch.cuda.synchronize()
torch.cuda.profiler.start()
it()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
G.check_overflow(dev)
print("ok")
