"""Binning utilities: mirror of gsplat/gsplat/utils.py:12-167 (same names, arguments, returned tensors).

These are the *staged, reference-typed* forms (int64 keys, int32 ids, host-side intersection count).  The operator
rasterize_gaussians_sum does not go through them on its fast path; it runs the fused sync-free chain
(csrc/binning.cu: gi2d_bin_fused).
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import cuda as _C


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                               cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int]) -> Tuple[Tensor, Tensor]:
    """utils.py:12-50 -> (isect_ids (n,) int64 = (tile<<32)|depth_bits, gaussian_ids (n,) int32)."""
    return _C.map_gaussian_to_intersects(num_points, num_intersects, xys.contiguous(), depths.contiguous(),
                                         radii.contiguous(), cum_tiles_hit.contiguous(), tile_bounds)


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor) -> Tensor:
    """utils.py:53-74 -> tile_bins (num_intersects, 2) int32, indexed by tile id (rows beyond are skipped, Q4)."""
    return _C.get_tile_bin_edges(num_intersects, isect_ids_sorted.contiguous())


def compute_cov2d_bounds(cov2d: Tensor) -> Tuple[Tensor, Tensor]:
    """utils.py:77-96 -> (conics (N,3), radii (N,1) float32)."""
    assert cov2d.shape[-1] == 3, (
        f"Expected input cov2d to be of shape (*batch, 3) (upper triangular values), but got {tuple(cov2d.shape)}"
    )
    num_pts = cov2d.shape[0]
    assert num_pts > 0
    return _C.compute_cov2d_bounds(num_pts, cov2d.contiguous())


def compute_cumulative_intersects(num_tiles_hit: Tensor) -> Tuple[int, Tensor]:
    """utils.py:99-118 -> (num_intersects as a Python int -- one device->host read, like `.item()` -- , cum int32)."""
    cum, total = _C.scan_tiles(num_tiles_hit.contiguous().to(torch.int32))
    return int(total.item()), cum


def _depths_known_zero(depths: Tensor) -> bool:
    return getattr(depths, "_gi2d_zero_depth_version", None) == depths._version


def bin_and_sort_gaussians(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                           cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int]
                           ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """utils.py:121-167 -> (isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins).

    The sort is stable (ties keep emission order = ascending gaussian id), which is what the reference gets from
    torch.sort on CUDA (quirk Q2).  tile_bins has max(num_intersects, tiles) rows so that every tile id is in range
    (the reference allocates num_intersects rows and indexes by tile id, quirk Q4).
    """
    isect_ids, gaussian_ids = map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit,
                                                         tile_bounds)
    n_tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    if _depths_known_zero(depths):
        begin, end = 32, 32 + max(1, (max(n_tiles, 2) - 1).bit_length())     # only the tile bits differ
    else:
        begin, end = 0, 64
    isect_ids_sorted, gaussian_ids_sorted = _C.sort_isects(isect_ids, gaussian_ids, begin, end)
    tile_bins = _C.get_tile_bin_edges(num_intersects, isect_ids_sorted, n_bins=max(num_intersects, n_tiles))
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins
