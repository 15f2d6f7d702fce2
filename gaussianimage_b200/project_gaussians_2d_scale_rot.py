"""project_gaussians_2d_scale_rot: (scale, rotation)-parametrised 2D Gaussians -> centre, conic, radius, tile count.

Mirror of the reference operator gsplat/gsplat/project_gaussians_2d_scale_rot.py:12-151 over csrc/project.cu.
"""
from __future__ import annotations

from typing import Tuple

from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
from ._lib import cpp_autograd
from .project_gaussians_2d import _tag_zero_depth


def project_gaussians_2d_scale_rot(
    means2d: Tensor,
    scales2d: Tensor,
    rotation: Tensor,
    img_height: int,
    img_width: int,
    tile_bounds: Tuple[int, int, int],
    clip_thresh: float = 0.01,
) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Returns (xys, depths, radii, conics, num_tiles_hit); differentiable w.r.t. means2d, scales2d, rotation."""
    T = cpp_autograd()
    if T is not None and means2d.shape[-2] >= 1 and means2d.shape[-1] == 2:     # else: the ValueError of the Python Function
        out = T.project_autograd(1, means2d.contiguous(), scales2d.contiguous(), rotation.contiguous(), img_height, img_width,
                                 tile_bounds[0], tile_bounds[1])
        _tag_zero_depth(out[1], out[0])
        return tuple(out)
    return _ProjectGaussians2dScaleRot.apply(
        means2d.contiguous(), scales2d.contiguous(), rotation.contiguous(), img_height, img_width, tile_bounds, clip_thresh
    )


class _ProjectGaussians2dScaleRot(Function):
    @staticmethod
    def forward(ctx, means2d, scales2d, rotation, img_height, img_width, tile_bounds, clip_thresh=0.01):
        num_points = means2d.shape[-2]
        if num_points < 1 or means2d.shape[-1] != 2:        # reference :76-77
            raise ValueError(f"Invalid shape for means2d: {means2d.shape}")
        xys, depths, radii, conics, num_tiles_hit = _C.project_gaussians_2d_scale_rot_forward(
            num_points, means2d, scales2d, rotation, img_height, img_width, tile_bounds, clip_thresh
        )
        ctx.img_height, ctx.img_width, ctx.num_points = img_height, img_width, num_points
        ctx.save_for_backward(means2d, scales2d, rotation, radii, conics)
        ctx.mark_non_differentiable(radii, num_tiles_hit)
        ctx.set_materialize_grads(False)
        _tag_zero_depth(depths, xys)
        return xys, depths, radii, conics, num_tiles_hit

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_num_tiles_hit):
        means2d, scales2d, rotation, radii, conics = ctx.saved_tensors
        v_mean2d, v_scale, v_rot = _C.project_backward_autograd("scale_rot", ctx.num_points, (scales2d, rotation), ctx.img_height,
                                                                ctx.img_width, radii, conics, v_xys, v_conics)
        return v_mean2d, v_scale, v_rot, None, None, None, None
