"""Alias package: `import gsplat` -> gaussianimage_b200 (2D render path only).

Put the repository root on PYTHONPATH and the reference's callers (`from gsplat.project_gaussians_2d import
project_gaussians_2d`, `from gsplat.rasterize_sum import rasterize_gaussians_sum`, `import gsplat.cuda as _C`, ...)
bind to the B200 kernels: the sub-module names of the reference are registered as the very same module objects.
The 3D path of the reference (project_gaussians, rasterize_gaussians, spherical_harmonics) is out of scope.
"""
import importlib as _il
import sys as _sys

from gaussianimage_b200 import *  # noqa: F401,F403
from gaussianimage_b200 import __all__, __version__  # noqa: F401

for _name in ("project_gaussians_2d", "project_gaussians_2d_scale_rot", "rasterize_sum", "utils", "cuda"):
    _mod = _il.import_module("gaussianimage_b200." + _name)
    _sys.modules[__name__ + "." + _name] = _mod
del _name, _mod
