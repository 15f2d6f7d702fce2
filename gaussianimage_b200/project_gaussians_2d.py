"""project_gaussians_2d: Cholesky-parametrised 2D Gaussians -> pixel centre, conic, radius, tile count.

Mirror of the reference operator gsplat/gsplat/project_gaussians_2d.py:12-141 (same name, arguments, return
tuple and gradient contract) over the sm_100a kernels of csrc/project.cu.
"""
from __future__ import annotations

from typing import Tuple

from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
from ._lib import cpp_autograd


_generation = 0


def _tag_zero_depth(depths: Tensor, xys: Tensor = None) -> None:
    """Both 2D projections write depths == 0 (foward2d.cu:67,122); remember that on the tensor object so that
    rasterize_gaussians_sum can sort on the tile bits alone (quirk Q2) without inspecting the data.  xys additionally gets
    the generation number of this projection call: the binning memo only ever matches tensors that carry the generation
    it was recorded for (never a buffer that merely has the same address and version counter)."""
    global _generation
    depths._gi2d_zero_depth_version = depths._version
    if xys is not None:
        _generation += 1
        xys._gi2d_gen = _generation


def project_gaussians_2d(
    means2d: Tensor,
    L_elements: Tensor,
    img_height: int,
    img_width: int,
    tile_bounds: Tuple[int, int, int],
    clip_thresh: float = 0.01,
) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Returns (xys (N,2), depths (N,), radii (N,) int32, conics (N,3), num_tiles_hit (N,) int32).

    Differentiable w.r.t. means2d and L_elements (through xys and conics); `clip_thresh` is ignored exactly as in
    the reference kernel (quirk Q11).
    """
    T = cpp_autograd()
    if T is not None:               # the same operator as a C++ autograd node (csrc/torch_glue.cpp: ProjectNode)
        out = T.project_autograd(0, means2d.contiguous(), L_elements.contiguous(), None, img_height, img_width,
                                 tile_bounds[0], tile_bounds[1])
        _tag_zero_depth(out[1], out[0])
        return tuple(out)
    return _ProjectGaussians2d.apply(
        means2d.contiguous(), L_elements.contiguous(), img_height, img_width, tile_bounds, clip_thresh
    )


class _ProjectGaussians2d(Function):
    @staticmethod
    def forward(ctx, means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh=0.01):
        num_points = means2d.shape[-2]
        xys, depths, radii, conics, num_tiles_hit = _C.project_gaussians_2d_forward(
            num_points, means2d, L_elements, img_height, img_width, tile_bounds, clip_thresh
        )
        ctx.img_height, ctx.img_width, ctx.num_points = img_height, img_width, num_points
        ctx.save_for_backward(means2d, L_elements, radii, conics)
        ctx.mark_non_differentiable(radii, num_tiles_hit)
        ctx.set_materialize_grads(False)      # v_depths etc. arrive as None instead of freshly filled zero tensors
        _tag_zero_depth(depths, xys)
        return xys, depths, radii, conics, num_tiles_hit

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_num_tiles_hit):
        means2d, L_elements, radii, conics = ctx.saved_tensors
        v_mean2d, v_L = _C.project_backward_autograd("cholesky", ctx.num_points, (L_elements,), ctx.img_height,
                                                     ctx.img_width, radii, conics, v_xys, v_conics)
        return v_mean2d, v_L, None, None, None, None
