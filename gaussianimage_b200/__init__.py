"""gaussianimage_b200 -- B200-native (sm_100a) implementation of GaussianImage's gsplat 2D render path.

Public surface = the reference's operator API for this path (gsplat/gsplat/__init__.py:1-16):
project_gaussians_2d, project_gaussians_2d_scale_rot, rasterize_gaussians_sum and the binning utilities,
all over the C ABI of libgi2d.so (include/gi2d.h).  `import gsplat` (the alias package at the repository root)
resolves to these modules so that the reference's gaussianimage_cholesky.py / gaussianimage_rs.py / train.py /
test_quantize.py run unchanged.  There is no CPU fallback.
"""
from .decode import decode_project_gaussians_2d
from .loss import rasterize_gaussians_sum_mse
from .project_gaussians_2d import project_gaussians_2d
from .project_gaussians_2d_scale_rot import project_gaussians_2d_scale_rot
from .rasterize_sum import check_overflow, rasterize_gaussians_sum, set_memoise_binning, set_sync_free
from .utils import (
    bin_and_sort_gaussians,
    compute_cov2d_bounds,
    compute_cumulative_intersects,
    get_tile_bin_edges,
    map_gaussian_to_intersects,
)

__version__ = "0.1.0"

__all__ = [
    "__version__",
    "project_gaussians_2d",
    "project_gaussians_2d_scale_rot",
    "rasterize_gaussians_sum",
    "bin_and_sort_gaussians",
    "compute_cumulative_intersects",
    "compute_cov2d_bounds",
    "get_tile_bin_edges",
    "map_gaussian_to_intersects",
    "set_sync_free",
    "check_overflow",
    "set_memoise_binning",
    "decode_project_gaussians_2d",
    "rasterize_gaussians_sum_mse",
]
