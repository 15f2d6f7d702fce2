"""Codec decode prologue fused into the Cholesky projection: the tensors of the reference's compressed
representation (compress_wo_ec, gaussianimage_cholesky.py:154-159) go straight into one kernel that does what
decompress_wo_ec does with seven torch kernels before it can rasterize (:161-166):

    means    = torch.tanh(xyz.float())
    cholesky = cholesky_quantizer.decompress(codes) + cholesky_bound          # codes * scale + beta, quantize.py:86-87
    colors   = features_dc_quantizer.decompress(index)                        # sum of codebook rows, quantize.py:146-150
    xys, depths, radii, conics, num_tiles_hit = project_gaussians_2d(means, cholesky, H, W, tile_bounds)

Forward only (the decode path runs under torch.no_grad(), test_quantize.py:68).  No CPU path.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import cuda as _C
from ._lib import check, lib
from .project_gaussians_2d import _tag_zero_depth


def _vec3(t: Tensor, name: str, dev) -> Tensor:
    t = t.detach()
    if t.device != dev or t.dtype != torch.float32 or t.numel() != 3:
        raise RuntimeError(f"{name} must be 3 float32 values on {dev}")
    return t.contiguous()


def decode_project_gaussians_2d(xyz: Tensor, quant_cholesky_elements: Tensor, scale: Tensor, beta: Tensor, cholesky_bound: Tensor,
                                feature_dc_index: Tensor, codebooks: Tensor, img_height: int, img_width: int,
                                tile_bounds: Tuple[int, int, int]):
    """xyz (N,2) fp16; quant_cholesky_elements (N,3) uint8 or float32 codes; scale / beta (3,) of the UniformQuantizer;
    cholesky_bound (1,3) or (3,); feature_dc_index (N,K) int32/int64; codebooks (K,C,3) fp32 (layer._codebook.embed[0]
    stacked).  Returns (xys, depths, radii, conics, num_tiles_hit, colors)."""
    if not xyz.is_cuda:
        raise RuntimeError("xyz must be a CUDA tensor")                      # bindings.h:9-14 convention
    dev = xyz.device
    if xyz.dtype != torch.float16 or xyz.dim() != 2 or xyz.size(1) != 2:
        raise RuntimeError("xyz must be (N,2) float16 (compress_wo_ec stores self._xyz.half())")
    N = xyz.size(0)
    codes = quant_cholesky_elements.detach()
    if codes.dtype not in (torch.uint8, torch.float32) or tuple(codes.shape) != (N, 3) or codes.device != dev:
        raise RuntimeError("quant_cholesky_elements must be (N,3) uint8 or float32 on the device of xyz")
    idx = feature_dc_index.detach()
    if idx.dtype not in (torch.int32, torch.int64) or idx.dim() != 2 or idx.size(0) != N or idx.device != dev:
        raise RuntimeError("feature_dc_index must be (N,K) int32 or int64 on the device of xyz")
    cb = codebooks.detach()
    if cb.dtype != torch.float32 or cb.dim() != 3 or cb.size(0) != idx.size(1) or cb.size(2) != 3 or cb.device != dev:
        raise RuntimeError("codebooks must be (K,C,3) float32 with K = feature_dc_index.shape[1]")
    xyz, codes, idx, cb = xyz.detach().contiguous(), codes.contiguous(), idx.contiguous(), cb.contiguous()
    scale, beta = _vec3(scale, "scale", dev), _vec3(beta, "beta", dev)
    bound = _vec3(cholesky_bound.reshape(-1), "cholesky_bound", dev)
    H, W = int(img_height), int(img_width)
    TX, TY = int(tile_bounds[0]), int(tile_bounds[1])
    with _C._guard(dev):
        xys = torch.empty((N, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((N,), dtype=torch.float32, device=dev)
        radii = torch.empty((N,), dtype=torch.int32, device=dev)
        conics = torch.empty((N, 3), dtype=torch.float32, device=dev)
        nth = torch.empty((N,), dtype=torch.int32, device=dev)
        colors = torch.empty((N, 3), dtype=torch.float32, device=dev)
        check(lib().gi2d_decode_project_cholesky_fwd(N, xyz.data_ptr(), codes.data_ptr(), int(codes.dtype == torch.float32),
                                                     scale.data_ptr(), beta.data_ptr(), bound.data_ptr(), idx.data_ptr(),
                                                     int(idx.dtype == torch.int64), idx.size(1), cb.size(1), cb.data_ptr(), H, W, TX, TY,
                                                     xys.data_ptr(), depths.data_ptr(), radii.data_ptr(), conics.data_ptr(),
                                                     nth.data_ptr(), colors.data_ptr(), _C._stream(dev)), "decode_project")
        _tag_zero_depth(depths, xys)        # rasterize_gaussians_sum may take the fused (tile-only sort) path
    return xys, depths, radii, conics, nth, colors
