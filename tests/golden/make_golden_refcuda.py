"""Generates tests/golden/ref_cuda_kernels.npz from the REFERENCE's own CUDA kernels, run on the GPU box.

The reference has no CPU implementation and no test for the 2D projection, the accumulated-sum rasterizer or their
backward passes, so these functions are pinned against outputs of the reference kernels
themselves: oracle/Makefile compiles foward2d.cu / backward2d.cu / forward.cu / backward.cu from where they lie under
/root/reference into oracle/_ref/libgsplat_ref_{jit,fast}.so (never copied), tests/refcuda.py launches them with the
reference's launch shapes, and this script stores inputs and outputs of a few small seeded scenes.  Staging between the
kernels (cumsum, torch.sort, gather) is the reference's Python (gsplat/gsplat/utils.py:116-167).

    gpurun -- 'python tests/golden/make_golden_refcuda.py gpurun_out/ref_cuda_kernels.npz'
    cp gpurun_out/ref_cuda_kernels.npz tests/golden/

`jit` = the reference's JIT build flags (-O3; gsplat/gsplat/cuda/_backend.py:39-40), the arithmetic the oracle and the
product restate bit for bit; `fast` = its pip build (-O3 --use_fast_math; gsplat/setup.py:74-86), stored for the
integer outputs of projection only, to document where the two reference builds differ (quirk Q8).
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path[:0] = [str(ROOT), str(ROOT / "tests")]
import refcuda  # noqa: E402
from gaussianimage_b200.synth import make_scene  # noqa: E402

SCENES = {
    "ragged": dict(H=70, W=100, N=257, seed=11),                 # image not a multiple of 16
    "small": dict(H=96, W=128, N=1500, seed=12),
    "dense": dict(H=64, W=96, N=4000, seed=13, dense=0.8),       # > 256 intersections in some tiles (quirk Q1)
}


def npf(t):
    return t.detach().cpu().numpy()


def one_scene(name, kw, out):
    s = make_scene(**kw)
    d = s.to("cuda:0")
    H, W = s.H, s.W
    TX, TY = s.tile_bounds[:2]
    g = torch.Generator().manual_seed(kw["seed"] + 100)
    v_out = (torch.randn(H, W, 3, generator=g) * 1e-2).to("cuda:0")
    pre = name + "_"
    for k in ("means2d", "cholesky", "scales", "rotation", "colors", "opacity"):
        out[pre + k] = npf(getattr(s, k))
    out[pre + "v_out"] = npf(v_out)
    out[pre + "HW"] = np.array([H, W])

    # ---- Cholesky chain: projection -> binning -> rasterize-sum fwd -> bwd -> projection bwd (jit build)
    xys, depths, radii, conics, nth = refcuda.project_cholesky_fwd(d.means2d, d.cholesky, H, W, "jit")
    n, cum, ids, gids, ids_s, gids_s, bins = refcuda.bin_and_sort(xys, depths, radii, nth, H, W, "jit")
    img, Ts, fidx = refcuda.rasterize_sum_fwd(H, W, gids_s, bins, xys, conics, d.colors, d.opacity, "jit")
    v_xy, v_conic, v_col, v_op = refcuda.rasterize_sum_bwd(H, W, gids_s, bins, xys, conics, d.colors, d.opacity, Ts, fidx,
                                                           v_out, "jit")
    v_cov, v_mean, v_L = refcuda.project_cholesky_bwd(d.means2d, d.cholesky, H, W, radii, conics, v_xy, v_conic, "jit")
    for k, v in dict(xys=xys, depths=depths, radii=radii, conics=conics, nth=nth, cum=cum, ids=ids, gids=gids, ids_sorted=ids_s,
                     gids_sorted=gids_s, bins=bins[:TX * TY], img=img, final_Ts=Ts, final_idx=fidx, v_xy=v_xy, v_conic=v_conic,
                     v_colors=v_col, v_opacity=v_op, v_cov2d=v_cov, v_mean2d=v_mean, v_L=v_L).items():
        out[pre + "chol_" + k] = npf(v)
    fx = refcuda.project_cholesky_fwd(d.means2d, d.cholesky, H, W, "fast")
    out[pre + "chol_fast_radii"], out[pre + "chol_fast_nth"] = npf(fx[2]), npf(fx[4])

    # ---- scale-rotation projection forward + backward (jit build), fed with the Cholesky chain's v_xy / v_conic
    rx = refcuda.project_scale_rot_fwd(d.means2d, d.scales, d.rotation, H, W, "jit")
    rb = refcuda.project_scale_rot_bwd(d.means2d, d.scales, d.rotation, H, W, rx[2], rx[3], v_xy, v_conic, "jit")
    for k, v in dict(xys=rx[0], radii=rx[2], conics=rx[3], nth=rx[4], v_cov2d=rb[0], v_mean2d=rb[1], v_scale=rb[2], v_rot=rb[3]).items():
        out[pre + "rs_" + k] = npf(v)


if __name__ == "__main__":
    assert torch.cuda.is_available() and refcuda.available("jit") and refcuda.available("fast")
    dst = Path(sys.argv[1] if len(sys.argv) > 1 else ROOT / "gpurun_out" / "ref_cuda_kernels.npz")
    data = {}
    for nm, kw in SCENES.items():
        one_scene(nm, kw, data)
    data["torch_version"] = np.array(torch.__version__)
    data["device"] = np.array(torch.cuda.get_device_name(0))
    dst.parent.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(dst, **data)
    print("wrote", dst, dst.stat().st_size, "bytes;", len(data), "arrays")
