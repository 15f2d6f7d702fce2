"""Generates tests/golden/ref_torch_impl.npz from the REFERENCE's own pure-PyTorch code (run in the build
container, where /root/reference exists; the GPU box only reads the committed .npz).

What is pinned: the three on-path functions the reference's test-suite covers
(gsplat/tests/test_cov2d_bounds.py:9-35, test_map_gaussians.py:9-73, test_get_tile_bin_edges.py:9-81), with the
same seeds and shapes those tests use, evaluated with gsplat/gsplat/_torch_impl.py on the CPU.

    python tests/golden/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, "/root/reference/gsplat")
from gsplat import _torch_impl  # noqa: E402  (the reference package; lazy CUDA loader, imports on CPU)

OUT = Path(__file__).resolve().parent / "ref_torch_impl.npz"


def cov2d_case():
    torch.manual_seed(42)
    _covs2d = torch.rand((100, 2, 2), dtype=torch.float32)
    covs2d = torch.stack([torch.triu(_covs2d)[:, 0, 0], torch.triu(_covs2d)[:, 0, 1], torch.triu(_covs2d)[:, 1, 1]], dim=-1)
    conic, radii, mask = _torch_impl.compute_cov2d_bounds(_covs2d)
    # NOTE: _torch_impl reads cov2d[...,0,1] of the full matrix == the triu value, so both see the same numbers
    return dict(cov_in=covs2d.numpy(), cov_conic=conic.numpy(), cov_radii=radii.numpy(), cov_mask=mask.numpy())


def map_case():
    torch.manual_seed(42)
    num_points = 100
    means3d = torch.randn((num_points, 3))
    scales = torch.randn((num_points, 3))
    quats = torch.randn((num_points, 4))
    quats /= torch.linalg.norm(quats, dim=-1, keepdim=True)
    viewmat, projmat = torch.eye(4), torch.eye(4)
    H, W = 512, 512
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)
    _, xys, depths, radii, conics, nth, masks = _torch_impl.project_gaussians_forward(
        means3d, scales, 0.3, quats, viewmat, projmat, 3.0, 3.0, (H, W), tile_bounds, 0.01)
    xys, depths, radii, nth = xys[masks], depths[masks].contiguous(), radii[masks], nth[masks]
    n_pts = int(masks.sum())
    cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
    n_isect = int(cum[-1])
    ids, gids = _torch_impl.map_gaussian_to_intersects(n_pts, xys, depths, radii, cum, tile_bounds)
    ids_sorted, order = torch.sort(ids, stable=True)
    gids_sorted = torch.gather(gids, 0, order)
    bins = _torch_impl.get_tile_bin_edges(n_isect, ids_sorted)
    return dict(map_xys=xys.numpy(), map_depths=depths.numpy(), map_radii=radii.to(torch.int32).numpy(),
                map_nth=nth.to(torch.int32).numpy(), map_cum=cum.numpy(), map_tile_bounds=np.array(tile_bounds),
                map_ids=ids.numpy(), map_gids=gids.numpy(), map_ids_sorted=ids_sorted.numpy(),
                map_gids_sorted=gids_sorted.numpy(), map_bins=bins.numpy())


if __name__ == "__main__":
    data = {}
    data.update(cov2d_case())
    data.update(map_case())
    np.savez_compressed(OUT, **data)
    print("wrote", OUT, {k: v.shape for k, v in data.items()})
