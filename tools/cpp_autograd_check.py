"""Python autograd Functions vs their C++ twins (GI2D_CPP_AUTOGRAD, csrc/torch_glue.cpp): bit-identical image and
gradients on a small scene (Cholesky and scale-rotation), then host time per eager fit step of each.  A few seconds.
The full check is the GPU suite under GI2D_CPP_AUTOGRAD=1."""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200 import _lib
from gaussianimage_b200.synth import make_scene

s = make_scene(128, 192, 3000, seed=1)
d = s.to("cuda:0")
rot = torch.rand(3000, 1, device="cuda:0") * 6.28
scales = d.cholesky[:, [0, 2]].contiguous()


def step(mode, leaves):
    if mode == "chol":
        proj = G.project_gaussians_2d(leaves[0], leaves[1], s.H, s.W, s.tile_bounds)
    else:
        proj = G.project_gaussians_2d_scale_rot(leaves[0], leaves[1], leaves[2], s.H, s.W, s.tile_bounds)
    out = G.rasterize_gaussians_sum(*proj, leaves[-1], d.opacity, s.H, s.W)
    loss = torch.nn.functional.mse_loss(out, d.gt)
    return out, torch.autograd.grad(loss, leaves)


res = {}
for mode in ("chol", "rs"):
    base = [d.means2d, d.cholesky, d.colors] if mode == "chol" else [d.means2d, scales, rot, d.colors]
    got = {}
    for cpp in (False, True):
        _lib.set_cpp_autograd(cpp)
        leaves = [t.clone().requires_grad_(True) for t in base]
        out, grads = step(mode, leaves)
        for _ in range(20):
            step(mode, leaves)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(300):
            step(mode, leaves)
        host = time.perf_counter() - t0
        torch.cuda.synchronize()
        got[cpp] = (out, grads, 1e6 * host / 300)
    # the image is deterministic; the gradients are summed with red.global.add across tiles (csrc/raster.cu), so they differ in
    # the last bits from run to run on EITHER path: report the Python path against itself as the yardstick
    _lib.set_cpp_autograd(False)
    again = step(mode, [t.clone().requires_grad_(True) for t in base])
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
    res[mode] = {"image_bit_identical": bool(torch.equal(got[False][0], got[True][0])),
                 "grad_max_rel_diff_cpp_vs_python": max(rel(a, b) for a, b in zip(got[True][1], got[False][1])),
                 "grad_max_rel_diff_python_vs_python": max(rel(a, b) for a, b in zip(again[1], got[False][1])),
                 "host_us_python_functions": got[False][2], "host_us_cpp_functions": got[True][2]}
print(json.dumps(res))
