"""Loss epilogue fused into the rasterizer.  The reference's training step turns the rendered image into a
loss with five more torch kernels and their autograd mirror images (gaussianimage_cholesky.py:84-85, :104; utils.py:20-25):

    image = torch.clamp(out_img, 0, 1).view(-1, H, W, 3).permute(0, 3, 1, 2).contiguous()
    loss  = F.mse_loss(image, gt_image)                     # loss_fn(..., 'L2')

`rasterize_gaussians_sum_mse` is rasterize_gaussians_sum whose forward kernel ends in that loss: per pixel it clamps,
subtracts the target (read in the caller's layout, (1,3,H,W) or (H,W,3)), accumulates the squared error and stores
d loss / d out_img directly into the buffer the backward rasterizer reads -- the image, the clamp, the permuted copy and
the loss gradient are never materialised.  The loss value is deterministic (per-CTA partial sums added in index order).
Callers of the reference must opt in (their loss lives in their own code); bench.py reports it as an extra line.
"""
from __future__ import annotations

import torch
from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
from ._lib import check, lib
from .rasterize_sum import _default_background, _forward_fused
from .utils import _depths_known_zero

_ws: dict = {}


def _workspace(dev):
    w = _ws.get(dev.index)
    if w is None:
        w = _ws[dev.index] = torch.zeros(lib().gi2d_raster_mse_workspace_bytes(), dtype=torch.uint8, device=dev)   # ticket = 0
    return w


class _RasterizeSumMSE(Function):
    @staticmethod
    def forward(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity, img_height, img_width, gt_image, clamp,
                background, return_image):
        H, W = int(img_height), int(img_width)
        if colors.shape[-1] != 3:
            raise AttributeError("only 3-channel colors are supported")
        if not (xys.is_cuda and xys.dtype == torch.float32 and _depths_known_zero(depths)):
            raise RuntimeError("rasterize_gaussians_sum_mse needs the outputs of project_gaussians_2d[_scale_rot] on a CUDA device")
        gt = gt_image.detach()
        if gt.dtype != torch.float32 or gt.device != xys.device or not gt.is_contiguous():
            raise RuntimeError("gt_image must be a contiguous float32 tensor on the device of xys")
        if tuple(gt.shape) in ((1, 3, H, W), (3, H, W)):
            chw = 1
        elif tuple(gt.shape) in ((H, W, 3), (1, H, W, 3)):
            chw = 0
        else:
            raise ValueError(f"gt_image must be (1,3,{H},{W}) or ({H},{W},3), got {tuple(gt.shape)}")
        TX, TY = (W + 15) // 16, (H + 15) // 16
        ctx.H, ctx.W, ctx.TX, ctx.TY, ctx.N = H, W, TX, TY, xys.size(0)
        ctx.set_materialize_grads(False)
        dev = xys.device
        L = lib()
        ws = _workspace(dev)

        def raster(r, status_ptr, stream):
            v_out = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
            image = torch.empty((H, W, 3), dtype=torch.float32, device=dev) if return_image else None
            loss = torch.empty((), dtype=torch.float32, device=dev)
            check(L.gi2d_raster_sum_fwd_packed_mse(TX, TY, W, H, r.tile_bins.data_ptr(), r.packed.data_ptr(), status_ptr,
                                                   bg.data_ptr(), gt.data_ptr(), chw, int(bool(clamp)), v_out.data_ptr(),
                                                   _C._p(image), loss.data_ptr(), ws.data_ptr(), ws.numel(), stream),
                  "rasterize_sum_forward_mse")
            r.extra = (v_out, image, loss)

        bg = _default_background(colors) if background is None else _C._f32(background.to(dev).contiguous(), "background")
        # the (N,12) gradient rows of the backward pass are allocated here and zero-filled by the binning chain (which touches
        # every Gaussian anyway): no memset in front of the backward rasterizer
        want_grads = any(ctx.needs_input_grad[k] for k in (0, 3, 5, 6))          # xys, conics, colors, opacity
        grads = torch.empty((xys.size(0), 12), dtype=torch.float32, device=dev) if want_grads else None
        r = _forward_fused(ctx, xys, radii, conics, num_tiles_hit, colors, opacity, bg, False, raster, zero_g12=grads)
        v_out, image, loss = r.extra
        ctx.fresh_grads = grads
        ctx.save_for_backward(r.tile_bins, r.packed, v_out)
        if return_image:
            ctx.mark_non_differentiable(image)
            return loss, image
        return loss

    @staticmethod
    def backward(ctx, v_loss, v_image=None):
        if v_loss is None:
            return (None,) * 13
        tile_bins, packed, v_out = ctx.saved_tensors
        N, H, W, TX, TY = ctx.N, ctx.H, ctx.W, ctx.TX, ctx.TY
        dev = packed.device
        scale = _C._f32(v_loss.contiguous(), "v_loss")
        with _C._guard(dev):
            grads, ctx.fresh_grads = getattr(ctx, "fresh_grads", None), None       # zeroed by the forward's chain: one use only
            if grads is not None:
                check(lib().gi2d_raster_sum_bwd_packed_acc(N, TX, TY, W, H, tile_bins.data_ptr(), packed.data_ptr(), v_out.data_ptr(),
                                                           scale.data_ptr(), grads.data_ptr(), _C._stream(dev)),
                      "rasterize_sum_backward")
            else:                                   # a second backward through the same node: rows allocated and zeroed here
                grads = torch.empty((N, 12), dtype=torch.float32, device=dev)
                check(lib().gi2d_raster_sum_bwd_packed_scaled(N, TX, TY, W, H, tile_bins.data_ptr(), packed.data_ptr(), v_out.data_ptr(),
                                                              scale.data_ptr(), grads.data_ptr(), _C._stream(dev)),
                      "rasterize_sum_backward")
        v_opacity = grads[:, 8:9].reshape(ctx.opacity_shape)
        return (grads[:, 0:2], None, None, grads[:, 2:5], None, grads[:, 5:8], v_opacity, None, None, None, None, None, None)


def rasterize_gaussians_sum_mse(xys: Tensor, depths: Tensor, radii: Tensor, conics: Tensor, num_tiles_hit: Tensor, colors: Tensor,
                                opacity: Tensor, img_height: int, img_width: int, gt_image: Tensor, clamp: bool = True,
                                background: Tensor = None, return_image: bool = False):
    """loss = F.mse_loss(clamp(rasterize_gaussians_sum(...), 0, 1) in the layout of gt_image, gt_image), differentiable w.r.t.
    xys / conics / colors / opacity exactly like rasterize_gaussians_sum followed by those torch ops.  gt_image is
    (1,3,H,W) (the reference's layout) or (H,W,3).  return_image=True also returns the clamped (H,W,3) image (no gradient).
    clamp=False corresponds to the models' `no_clamp` option."""
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255                       # rasterize_sum.py:53-55
    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise ValueError("xys must have dimensions (N, 2)")  # rasterize_sum.py:66-67
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    return _RasterizeSumMSE.apply(xys.contiguous(), depths.contiguous(), radii.contiguous(), conics.contiguous(),
                                  num_tiles_hit.contiguous(), colors.contiguous(), opacity.contiguous(), img_height, img_width,
                                  gt_image, clamp, background, return_image)
