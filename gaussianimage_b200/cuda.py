"""Binding layer: the same function names, argument order and return tuples as the reference's pybind11 module
(gsplat/gsplat/cuda/csrc/ext.cpp:10-23, re-exported lazily by gsplat/gsplat/cuda/__init__.py:14-30), implemented
over the C ABI of libgi2d.so.  Tensors in, tensors out; allocation happens here (torch.empty -- the kernels write
every element), the current torch stream is used (the reference uses the legacy default stream, quirk Q9).

Error behaviour follows bindings.h:9-14 (CHECK_INPUT): a CPU or non-contiguous tensor raises RuntimeError.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from ._lib import check, glue, lib


def _chk(x: Tensor, name: str) -> Tensor:
    if not x.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not x.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    return x


def _f32(x: Tensor, name: str) -> Tensor:
    _chk(x, name)
    if x.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {x.dtype}")
    return x


def _i32(x: Tensor, name: str) -> Tensor:
    _chk(x, name)
    if x.dtype != torch.int32:
        raise RuntimeError(f"{name} must be int32, got {x.dtype}")
    return x


class _NoGuard:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NOGUARD = _NoGuard()


def _guard(device):
    """Device guard that costs nothing when `device` is already current (the usual case; torch.cuda.device() costs
    several microseconds per use, which matters on a path whose whole forward is ~60 us of GPU time)."""
    if device.index is None or device.index == torch._C._cuda_getDevice():
        return _NOGUARD
    return torch.cuda.device(device)


def _stream(device) -> int:
    """Raw cudaStream_t of torch's current stream on `device` (the reference uses the legacy default stream, Q9)."""
    idx = device.index
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice() if idx is None else idx)


def _p(x) -> int:
    return 0 if x is None else x.data_ptr()


def project_gaussians_2d_forward(num_points: int, means2d: Tensor, L_elements: Tensor, img_height: int,
                                 img_width: int, tile_bounds: Tuple[int, int, int], clip_thresh: float = 0.01):
    """ext.cpp:16 -> bindings.cu:781-839. Returns (xys, depths, radii, conics, num_tiles_hit)."""
    T = glue()
    if T is not None:           # same checks, allocations and launch as below, as one C++ call (csrc/torch_glue.cpp)
        return T.project_fwd(0, int(num_points), means2d, L_elements, None, img_height, img_width, tile_bounds[0], tile_bounds[1])
    means2d, L = _f32(means2d, "means2d"), _f32(L_elements, "L_elements")
    N, dev = int(num_points), means2d.device
    with _guard(dev):
        xys = torch.empty((N, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((N,), dtype=torch.float32, device=dev)
        radii = torch.empty((N,), dtype=torch.int32, device=dev)
        conics = torch.empty((N, 3), dtype=torch.float32, device=dev)
        nth = torch.empty((N,), dtype=torch.int32, device=dev)
        check(lib().gi2d_project_cholesky_fwd(N, _p(means2d), _p(L), img_height, img_width, tile_bounds[0], tile_bounds[1],
                                              _p(xys), _p(depths), _p(radii), _p(conics), _p(nth), _stream(dev)),
              "project_gaussians_2d_forward")
    return xys, depths, radii, conics, nth


def project_gaussians_2d_scale_rot_forward(num_points: int, means2d: Tensor, scales2d: Tensor, rotation: Tensor,
                                           img_height: int, img_width: int, tile_bounds: Tuple[int, int, int],
                                           clip_thresh: float = 0.01):
    """ext.cpp:18 -> bindings.cu:841-899."""
    T = glue()
    if T is not None:
        return T.project_fwd(1, int(num_points), means2d, scales2d, rotation, img_height, img_width, tile_bounds[0],
                             tile_bounds[1])
    means2d, scales2d, rotation = _f32(means2d, "means2d"), _f32(scales2d, "scales2d"), _f32(rotation, "rotation")
    N, dev = int(num_points), means2d.device
    with _guard(dev):
        xys = torch.empty((N, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((N,), dtype=torch.float32, device=dev)
        radii = torch.empty((N,), dtype=torch.int32, device=dev)
        conics = torch.empty((N, 3), dtype=torch.float32, device=dev)
        nth = torch.empty((N,), dtype=torch.int32, device=dev)
        check(lib().gi2d_project_scale_rot_fwd(N, _p(means2d), _p(scales2d), _p(rotation), img_height, img_width,
                                               tile_bounds[0], tile_bounds[1], _p(xys), _p(depths), _p(radii),
                                               _p(conics), _p(nth), _stream(dev)),
              "project_gaussians_2d_scale_rot_forward")
    return xys, depths, radii, conics, nth


def project_gaussians_2d_backward(num_points: int, means2d: Tensor, L_elements: Tensor, img_height: int,
                                  img_width: int, radii: Tensor, conics: Tensor, v_xy: Tensor, v_depth: Tensor,
                                  v_conic: Tensor, want_cov2d: bool = True):
    """ext.cpp:17 -> bindings.cu:902-949. Returns (v_cov2d, v_mean2d, v_L_elements)."""
    L, radii, conics = _f32(L_elements, "L_elements"), _i32(radii, "radii"), _f32(conics, "conics")
    v_xy, v_conic = _f32(v_xy.contiguous(), "v_xy"), _f32(v_conic.contiguous(), "v_conic")
    N, dev = int(num_points), L.device
    with _guard(dev):
        v_cov2d = torch.empty((N, 3), dtype=torch.float32, device=dev) if want_cov2d else None
        v_mean2d = torch.empty((N, 2), dtype=torch.float32, device=dev)
        v_L = torch.empty((N, 3), dtype=torch.float32, device=dev)
        check(lib().gi2d_project_cholesky_bwd(N, _p(L), img_height, img_width, _p(radii), _p(conics), _p(v_xy), _p(v_conic),
                                              _p(v_cov2d), _p(v_mean2d), _p(v_L), _stream(dev)),
              "project_gaussians_2d_backward")
    return v_cov2d, v_mean2d, v_L


def project_gaussians_2d_scale_rot_backward(num_points: int, means2d: Tensor, scales2d: Tensor, rotation: Tensor,
                                            img_height: int, img_width: int, radii: Tensor, conics: Tensor,
                                            v_xy: Tensor, v_depth: Tensor, v_conic: Tensor, want_cov2d: bool = True):
    """ext.cpp:19 -> bindings.cu:951-1004. Returns (v_cov2d, v_mean2d, v_scale (N,2), v_rot (N,1))."""
    scales2d, rotation = _f32(scales2d, "scales2d"), _f32(rotation, "rotation")
    radii, conics = _i32(radii, "radii"), _f32(conics, "conics")
    v_xy, v_conic = _f32(v_xy.contiguous(), "v_xy"), _f32(v_conic.contiguous(), "v_conic")
    N, dev = int(num_points), scales2d.device
    with _guard(dev):
        v_cov2d = torch.empty((N, 3), dtype=torch.float32, device=dev) if want_cov2d else None
        v_mean2d = torch.empty((N, 2), dtype=torch.float32, device=dev)
        v_scale = torch.empty((N, 2), dtype=torch.float32, device=dev)
        v_rot = torch.empty((N, 1), dtype=torch.float32, device=dev)
        check(lib().gi2d_project_scale_rot_bwd(N, _p(scales2d), _p(rotation), img_height, img_width, _p(radii), _p(conics),
                                               _p(v_xy), _p(v_conic), _p(v_cov2d), _p(v_mean2d), _p(v_scale), _p(v_rot),
                                               _stream(dev)),
              "project_gaussians_2d_scale_rot_backward")
    return v_cov2d, v_mean2d, v_scale, v_rot


def compute_cov2d_bounds(num_pts: int, cov2d: Tensor):
    """ext.cpp:12 -> bindings.cu:41-61. Returns (conics (N,3), radii (N,1) float32)."""
    cov2d = _f32(cov2d, "cov2d")
    N, dev = int(num_pts), cov2d.device
    with _guard(dev):
        conics = torch.empty((N, 3), dtype=torch.float32, device=dev)
        radii = torch.empty((N, 1), dtype=torch.float32, device=dev)
        check(lib().gi2d_cov2d_bounds(N, _p(cov2d), _p(conics), _p(radii), _stream(dev)), "compute_cov2d_bounds")
    return conics, radii


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                               cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int]):
    """ext.cpp:13 -> bindings.cu:274-313. Returns (isect_ids int64, gaussian_ids int32)."""
    xys, depths = _f32(xys, "xys"), _f32(depths, "depths")
    radii, cum = _i32(radii, "radii"), _i32(cum_tiles_hit, "cum_tiles_hit")
    dev = xys.device
    with _guard(dev):
        ids = torch.empty((num_intersects,), dtype=torch.int64, device=dev)
        gids = torch.empty((num_intersects,), dtype=torch.int32, device=dev)
        check(lib().gi2d_emit_isects(int(num_points), _p(xys), _p(depths), _p(radii), _p(cum), tile_bounds[0], tile_bounds[1],
                                     int(num_intersects), _p(ids), _p(gids), _stream(dev)), "map_gaussian_to_intersects")
    return ids, gids


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor, n_bins: int | None = None):
    """ext.cpp:14 -> bindings.cu:315-330. tile_bins (num_intersects,2) like the reference unless n_bins is given."""
    _chk(isect_ids_sorted, "isect_ids_sorted")
    if isect_ids_sorted.dtype != torch.int64:
        raise RuntimeError("isect_ids_sorted must be int64")
    dev = isect_ids_sorted.device
    rows = int(num_intersects) if n_bins is None else int(n_bins)
    with _guard(dev):
        bins = torch.empty((rows, 2), dtype=torch.int32, device=dev)
        check(lib().gi2d_tile_bins(int(num_intersects), _p(isect_ids_sorted), rows, _p(bins), _stream(dev)), "get_tile_bin_edges")
    return bins


def sort_isects(isect_ids: Tensor, gaussian_ids: Tensor, begin_bit: int = 0, end_bit: int = 64):
    """torch.sort + torch.gather of utils.py:164-165 as one stable onesweep radix sort."""
    _chk(isect_ids, "isect_ids"); _i32(gaussian_ids, "gaussian_ids")
    n, dev = isect_ids.numel(), isect_ids.device
    with _guard(dev):
        ids_s, gids_s = torch.empty_like(isect_ids), torch.empty_like(gaussian_ids)
        if n:
            nbytes = lib().gi2d_sort_workspace_bytes(n, begin_bit, end_bit)
            ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            check(lib().gi2d_sort_isects(n, _p(isect_ids), _p(gaussian_ids), _p(ids_s), _p(gids_s), begin_bit, end_bit, _p(ws),
                                         nbytes, _stream(dev)), "sort_isects")
    return ids_s, gids_s


def scan_tiles(num_tiles_hit: Tensor):
    """torch.cumsum(int32) of utils.py:116 -> (cum (N,) int32, total (1,) int32 device tensor)."""
    nth = _i32(num_tiles_hit, "num_tiles_hit")
    N, dev = nth.numel(), nth.device
    with _guard(dev):
        cum = torch.empty_like(nth)
        total = torch.empty((1,), dtype=torch.int32, device=dev)
        nbytes = lib().gi2d_scan_workspace_bytes(N)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
        check(lib().gi2d_scan_tiles(N, _p(nth), _p(cum), _p(total), _p(ws), nbytes, _stream(dev)), "scan_tiles")
    return cum, total


def rasterize_sum_forward(tile_bounds, block, img_size, gaussian_ids_sorted: Tensor, tile_bins: Tensor, xys: Tensor,
                          conics: Tensor, colors: Tensor, opacities: Tensor, background: Tensor):
    """ext.cpp:22 -> bindings.cu:400-469. Returns (out_img (H,W,3), final_Ts (H,W), final_idx (H,W) int32)."""
    gids, bins = _i32(gaussian_ids_sorted, "gaussian_ids_sorted"), _i32(tile_bins, "tile_bins")
    xys, conics, colors, opac = _f32(xys, "xys"), _f32(conics, "conics"), _f32(colors, "colors"), _f32(opacities, "opacities")
    _chk(background, "background")
    if tuple(block[:2]) != (16, 16):
        raise ValueError("only 16x16 tiles are coherent in the reference (helpers.cuh:36-41) and supported here")
    W, H = int(img_size[0]), int(img_size[1])
    TX, TY = int(tile_bounds[0]), int(tile_bounds[1])
    if bins.shape[0] < TX * TY:
        raise RuntimeError(f"tile_bins has {bins.shape[0]} rows but the image has {TX * TY} tiles "
                           "(the reference reads out of bounds here, quirk Q4)")
    dev = xys.device
    with _guard(dev):
        out = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
        Ts = torch.empty((H, W), dtype=torch.float32, device=dev)
        fidx = torch.empty((H, W), dtype=torch.int32, device=dev)
        check(lib().gi2d_raster_sum_fwd(TX, TY, W, H, _p(gids), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac), _p(out),
                                        _p(Ts), _p(fidx), _stream(dev)), "rasterize_sum_forward")
    return out, Ts, fidx


def rasterize_sum_backward(img_height: int, img_width: int, BLOCK_H: int, BLOCK_W: int, gaussian_ids_sorted: Tensor,
                           tile_bins: Tensor, xys: Tensor, conics: Tensor, colors: Tensor, opacities: Tensor,
                           background: Tensor, final_Ts: Tensor, final_idx: Tensor, v_output: Tensor,
                           v_output_alpha: Tensor):
    """ext.cpp:23 -> bindings.cu:706-779. Returns (v_xy, v_conic, v_colors, v_opacity (N,1))."""
    gids, bins = _i32(gaussian_ids_sorted, "gaussian_ids_sorted"), _i32(tile_bins, "tile_bins")
    xys, conics, colors, opac = _f32(xys, "xys"), _f32(conics, "conics"), _f32(colors, "colors"), _f32(opacities, "opacities")
    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise RuntimeError("xys must have dimensions (num_points, 2)")          # bindings.cu:734-736
    if colors.ndimension() != 2 or colors.size(1) != 3:
        raise RuntimeError("colors must have 2 dimensions")                     # bindings.cu:738-740
    if (BLOCK_H, BLOCK_W) != (16, 16):
        raise ValueError("only 16x16 tiles are supported")
    v_out = _f32(v_output.contiguous(), "v_output")
    H, W = int(img_height), int(img_width)
    TX, TY = (W + 15) // 16, (H + 15) // 16
    N, dev = xys.size(0), xys.device
    with _guard(dev):
        v_xy = torch.empty((N, 2), dtype=torch.float32, device=dev)
        v_conic = torch.empty((N, 3), dtype=torch.float32, device=dev)
        v_colors = torch.empty((N, 3), dtype=torch.float32, device=dev)
        v_opacity = torch.empty((N, 1), dtype=torch.float32, device=dev)
        check(lib().gi2d_raster_sum_bwd(N, TX, TY, W, H, _p(gids), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac),
                                        _p(final_idx), _p(v_out), _p(v_xy), _p(v_conic), _p(v_colors), _p(v_opacity),
                                        _stream(dev)), "rasterize_sum_backward")
    return v_xy, v_conic, v_colors, v_opacity


def _is_packed_pair(v_xy: Tensor, v_conic: Tensor) -> bool:
    """True when v_xy (N,2) and v_conic (N,3) are columns 0-1 and 2-4 of one (N,12) gradient buffer written by
    gi2d_raster_sum_bwd_packed, so the projection backward can read the rows in place."""
    return (v_xy.is_cuda and v_xy.dtype == torch.float32 and v_conic.dtype == torch.float32
            and v_xy.dim() == 2 and v_conic.dim() == 2 and v_xy.shape[1] == 2 and v_conic.shape[1] == 3
            and v_xy.stride() == (12, 1) and v_conic.stride() == (12, 1)
            and v_conic.data_ptr() == v_xy.data_ptr() + 8 and v_xy.data_ptr() % 16 == 0)


def project_backward_autograd(mode: str, num_points: int, params, img_height: int, img_width: int, radii: Tensor,
                              conics: Tensor, v_xy, v_conic):
    """Backward of either projection for the autograd Functions: (v_mean2d, v_L) or (v_mean2d, v_scale, v_rot).
    Skips the v_cov2d output the reference computes and discards, and reads packed gradient rows in place."""
    N, dev = int(num_points), conics.device
    if v_xy is None:
        v_xy = torch.zeros((N, 2), dtype=torch.float32, device=dev)
    if v_conic is None:
        v_conic = torch.zeros((N, 3), dtype=torch.float32, device=dev)
    if not _is_packed_pair(v_xy, v_conic):
        if mode == "cholesky":
            return project_gaussians_2d_backward(N, None, params[0], img_height, img_width, radii, conics, v_xy, None, v_conic,
                                                 want_cov2d=False)[1:]
        return project_gaussians_2d_scale_rot_backward(N, None, params[0], params[1], img_height, img_width, radii, conics,
                                                       v_xy, None, v_conic, want_cov2d=False)[1:]
    T = glue()
    if T is not None:
        if mode == "cholesky":
            return T.project_bwd_packed(0, N, params[0], None, img_height, img_width, radii, conics, v_xy)[:2]
        return T.project_bwd_packed(1, N, params[0], params[1], img_height, img_width, radii, conics, v_xy)
    with _guard(dev):
        v_mean2d = torch.empty((N, 2), dtype=torch.float32, device=dev)
        if mode == "cholesky":
            v_L = torch.empty((N, 3), dtype=torch.float32, device=dev)
            check(lib().gi2d_project_cholesky_bwd_packed(N, _p(params[0]), img_height, img_width, _p(radii), _p(conics),
                                                         v_xy.data_ptr(), _p(v_mean2d), _p(v_L), _stream(dev)),
                  "project_gaussians_2d_backward")
            return v_mean2d, v_L
        v_scale = torch.empty((N, 2), dtype=torch.float32, device=dev)
        v_rot = torch.empty((N, 1), dtype=torch.float32, device=dev)
        check(lib().gi2d_project_scale_rot_bwd_packed(N, _p(params[0]), _p(params[1]), img_height, img_width, _p(radii),
                                                      _p(conics), v_xy.data_ptr(), _p(v_mean2d), _p(v_scale), _p(v_rot),
                                                      _stream(dev)), "project_gaussians_2d_scale_rot_backward")
        return v_mean2d, v_scale, v_rot
