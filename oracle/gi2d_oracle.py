"""numpy/ctypes front of oracle/gi2d_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may import
this module.  Function names follow the reference operators they restate (file:line in the C file).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "libgi2d_oracle.so"
_lib = None

_f = np.float32
_i = np.int32


def build(force: bool = False) -> Path:
    """Compile the C restatement (gcc) -- and, where /root/reference exists, the reference-kernel harness."""
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < (_HERE / "gi2d_oracle.c").stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "oracle"], check=True, capture_output=True)
    return _LIB_PATH


def build_ref() -> None:
    subprocess.run(["make", "-C", str(_HERE), "ref", "-j8"], check=True, capture_output=True)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(_LIB_PATH))
        _lib.gi2d_oracle_cumsum.restype = C.c_int32
        _lib.gi2d_oracle_count_evals.restype = C.c_int64
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def tile_bounds(H: int, W: int):
    return (W + 15) // 16, (H + 15) // 16


def project_gaussians_2d(means2d, L, H, W, tile_bounds_=None):
    """foward2d.cu:12-69 -> (xys, depths, radii, conics, num_tiles_hit)."""
    means2d, L = _c(means2d, _f), _c(L, _f)
    N = means2d.shape[0]
    TX, TY = (int(tile_bounds_[0]), int(tile_bounds_[1])) if tile_bounds_ is not None else tile_bounds(H, W)
    xys, depths = np.zeros((N, 2), _f), np.zeros(N, _f)
    radii, conics, nth = np.zeros(N, _i), np.zeros((N, 3), _f), np.zeros(N, _i)
    lib().gi2d_oracle_project_cholesky_fwd(N, _p(means2d), _p(L), H, W, TX, TY, _p(xys), _p(depths), _p(radii),
                                           _p(conics), _p(nth))
    return xys, depths, radii, conics, nth


def project_gaussians_2d_scale_rot(means2d, scales, rot, H, W, tile_bounds_=None):
    """foward2d.cu:71-123."""
    means2d, scales, rot = _c(means2d, _f), _c(scales, _f), _c(rot, _f).reshape(-1)
    N = means2d.shape[0]
    TX, TY = (int(tile_bounds_[0]), int(tile_bounds_[1])) if tile_bounds_ is not None else tile_bounds(H, W)
    xys, depths = np.zeros((N, 2), _f), np.zeros(N, _f)
    radii, conics, nth = np.zeros(N, _i), np.zeros((N, 3), _f), np.zeros(N, _i)
    lib().gi2d_oracle_project_scale_rot_fwd(N, _p(means2d), _p(scales), _p(rot), H, W, TX, TY, _p(xys), _p(depths),
                                            _p(radii), _p(conics), _p(nth))
    return xys, depths, radii, conics, nth


def compute_cov2d_bounds(cov2d):
    """bindings.cu:21-61 -> (conics (N,3), radii (N,1) float32)."""
    cov2d = _c(cov2d, _f)
    N = cov2d.shape[0]
    conics, radii = np.zeros((N, 3), _f), np.zeros((N, 1), _f)
    lib().gi2d_oracle_cov2d_bounds(N, _p(cov2d), _p(conics), _p(radii))
    return conics, radii


def compute_cumulative_intersects(num_tiles_hit):
    """utils.py:99-118 -> (int, cum int32)."""
    nth = _c(num_tiles_hit, _i)
    cum = np.zeros_like(nth)
    total = lib().gi2d_oracle_cumsum(nth.shape[0], _p(nth), _p(cum))
    return int(total), cum


def map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds_):
    """forward.cu:100-136 -> (isect_ids int64, gaussian_ids int32)."""
    xys, depths, radii, cum = _c(xys, _f), _c(depths, _f), _c(radii, _i), _c(cum_tiles_hit, _i)
    num_intersects = int(num_intersects)
    ids, gids = np.zeros(num_intersects, np.int64), np.zeros(num_intersects, _i)
    lib().gi2d_oracle_map_intersects(int(num_points), _p(xys), _p(depths), _p(radii), _p(cum), int(tile_bounds_[0]),
                                     int(tile_bounds_[1]), _p(ids), _p(gids))
    return ids, gids


def sort_intersects(isect_ids, gaussian_ids):
    """utils.py:164-165 (torch.sort + gather), stable."""
    ids, gids = _c(isect_ids, np.int64), _c(gaussian_ids, _i)
    ids_s, gids_s = np.zeros_like(ids), np.zeros_like(gids)
    lib().gi2d_oracle_sort_intersects(ids.shape[0], _p(ids), _p(gids), _p(ids_s), _p(gids_s))
    return ids_s, gids_s


def get_tile_bin_edges(num_intersects, isect_ids_sorted, n_bins=None):
    """forward.cu:141-163; n_bins rows (default num_intersects like bindings.cu:319-321)."""
    ids = _c(isect_ids_sorted, np.int64)
    num_intersects = int(num_intersects)
    n_bins = num_intersects if n_bins is None else int(n_bins)
    bins = np.zeros((n_bins, 2), _i)
    lib().gi2d_oracle_tile_bin_edges(num_intersects, _p(ids), n_bins, _p(bins))
    return bins


def bin_and_sort_gaussians(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds_):
    """utils.py:121-167; tile_bins gets max(num_intersects, TX*TY) rows (Q4 made safe)."""
    ids, gids = map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds_)
    ids_s, gids_s = sort_intersects(ids, gids)
    bins = get_tile_bin_edges(num_intersects, ids_s, max(int(num_intersects), int(tile_bounds_[0]) * int(tile_bounds_[1])))
    return ids, gids, ids_s, gids_s, bins


def rasterize_sum_forward(H, W, gids_sorted, tile_bins, xys, conics, colors, opac, want_ambig=False):
    """forward.cu:512-627 -> (out_img (H,W,3), final_Ts (H,W), final_idx (H,W) int32[, ambig (H,W) uint8])."""
    TX, TY = tile_bounds(H, W)
    gids, bins = _c(gids_sorted, _i), _c(tile_bins, _i)
    assert bins.shape[0] >= TX * TY, "tile_bins must cover every tile"
    xys, conics, colors, opac = _c(xys, _f), _c(conics, _f), _c(colors, _f), _c(opac, _f).reshape(-1)
    out, Ts, fidx = np.zeros((H, W, 3), _f), np.zeros((H, W), _f), np.zeros((H, W), _i)
    amb = np.zeros((H, W), np.uint8)
    lib().gi2d_oracle_raster_sum_fwd(TX, TY, W, H, _p(gids), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac),
                                     _p(out), _p(Ts), _p(fidx), _p(amb))
    return (out, Ts, fidx, amb) if want_ambig else (out, Ts, fidx)


def rasterize_sum_backward(H, W, gids_sorted, tile_bins, xys, conics, colors, opac, final_idx, v_out):
    """backward.cu:696-862 -> (v_xy, v_conic, v_colors, v_opacity (N,1))."""
    TX, TY = tile_bounds(H, W)
    gids, bins = _c(gids_sorted, _i), _c(tile_bins, _i)
    xys, conics, colors, opac = _c(xys, _f), _c(conics, _f), _c(colors, _f), _c(opac, _f).reshape(-1)
    fidx, v_out = _c(final_idx, _i), _c(v_out, _f)
    N = xys.shape[0]
    v_xy, v_conic, v_col, v_op = np.zeros((N, 2), _f), np.zeros((N, 3), _f), np.zeros((N, 3), _f), np.zeros((N, 1), _f)
    lib().gi2d_oracle_raster_sum_bwd(TX, TY, W, H, _p(gids), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac),
                                     _p(fidx), _p(v_out), N, _p(v_xy), _p(v_conic), _p(v_col), _p(v_op))
    return v_xy, v_conic, v_col, v_op


def project_gaussians_2d_backward(L, H, W, radii, conics, v_xy, v_conic):
    """backward2d.cu:8-51 -> (v_cov2d, v_mean2d, v_L)."""
    L, radii, conics, v_xy, v_conic = _c(L, _f), _c(radii, _i), _c(conics, _f), _c(v_xy, _f), _c(v_conic, _f)
    N = L.shape[0]
    v_cov, v_mean, v_L = np.zeros((N, 3), _f), np.zeros((N, 2), _f), np.zeros((N, 3), _f)
    lib().gi2d_oracle_project_cholesky_bwd(N, _p(L), H, W, _p(radii), _p(conics), _p(v_xy), _p(v_conic), _p(v_cov),
                                           _p(v_mean), _p(v_L))
    return v_cov, v_mean, v_L


def project_gaussians_2d_scale_rot_backward(scales, rot, H, W, radii, conics, v_xy, v_conic):
    """backward2d.cu:53-109 -> (v_cov2d, v_mean2d, v_scale, v_rot (N,1))."""
    scales, rot = _c(scales, _f), _c(rot, _f).reshape(-1)
    radii, conics, v_xy, v_conic = _c(radii, _i), _c(conics, _f), _c(v_xy, _f), _c(v_conic, _f)
    N = scales.shape[0]
    v_cov, v_mean = np.zeros((N, 3), _f), np.zeros((N, 2), _f)
    v_scale, v_rot = np.zeros((N, 2), _f), np.zeros((N, 1), _f)
    lib().gi2d_oracle_project_scale_rot_bwd(N, _p(scales), _p(rot), H, W, _p(radii), _p(conics), _p(v_xy), _p(v_conic),
                                            _p(v_cov), _p(v_mean), _p(v_scale), _p(v_rot))
    return v_cov, v_mean, v_scale, v_rot


def count_evals(H, W, tile_bins) -> int:
    TX, TY = tile_bounds(H, W)
    bins = _c(tile_bins, _i)
    return int(lib().gi2d_oracle_count_evals(TX, TY, W, H, _p(bins)))


# ----------------------------------------------------------------------------------------------
# Whole-path helpers (what a caller such as gaussianimage_cholesky.py:76-85 does with the operators)
# ----------------------------------------------------------------------------------------------

def render_cholesky(means2d, L, colors, opac, H, W, want_all=False):
    """project -> cumsum -> bin_and_sort -> rasterize_sum_forward; returns out_img (H,W,3) (zeros-based sum)."""
    TX, TY = tile_bounds(H, W)
    xys, depths, radii, conics, nth = project_gaussians_2d(means2d, L, H, W)
    n, cum = compute_cumulative_intersects(nth)
    if n < 1:   # rasterize_sum.py:121-129: background (ones) image
        out = np.ones((H, W, 3), _f)
        return (out, None) if want_all else out
    ids, gids, ids_s, gids_s, bins = bin_and_sort_gaussians(means2d.shape[0], n, xys, depths, radii, cum, (TX, TY, 1))
    out, Ts, fidx = rasterize_sum_forward(H, W, gids_s, bins, xys, conics, colors, opac)
    if want_all:
        return out, dict(xys=xys, depths=depths, radii=radii, conics=conics, num_tiles_hit=nth, cum=cum, n=n,
                         isect_ids=ids, gaussian_ids=gids, isect_ids_sorted=ids_s, gaussian_ids_sorted=gids_s,
                         tile_bins=bins, final_idx=fidx, final_Ts=Ts)
    return out


def fit_step_cholesky(means2d, L, colors, opac, gt, H, W):
    """One fwd+bwd of the hot path under an MSE loss on the unclamped image: returns (loss, grads dict)."""
    out, st = render_cholesky(means2d, L, colors, opac, H, W, want_all=True)
    diff = out - gt
    loss = float(np.mean(diff.astype(np.float64) ** 2))
    v_out = (2.0 / diff.size) * diff
    v_xy, v_conic, v_col, v_op = rasterize_sum_backward(H, W, st["gaussian_ids_sorted"], st["tile_bins"], st["xys"],
                                                       st["conics"], colors, opac, st["final_idx"], v_out)
    _, v_mean, v_L = project_gaussians_2d_backward(L, H, W, st["radii"], st["conics"], v_xy, v_conic)
    return loss, dict(v_means2d=v_mean, v_L=v_L, v_colors=v_col, v_opacity=v_op, v_xy=v_xy, v_conic=v_conic)


def fit_step_scale_rot(means2d, scales, rot, colors, opac, gt, H, W):
    """The same step for the scale-rotation model (gaussianimage_rs.py:74-95)."""
    TX, TY = tile_bounds(H, W)
    xys, depths, radii, conics, nth = project_gaussians_2d_scale_rot(means2d, scales, rot, H, W)
    n, cum = compute_cumulative_intersects(nth)
    ids, gids, ids_s, gids_s, bins = bin_and_sort_gaussians(means2d.shape[0], n, xys, depths, radii, cum, (TX, TY, 1))
    out, Ts, fidx = rasterize_sum_forward(H, W, gids_s, bins, xys, conics, colors, opac)
    diff = out - gt
    loss = float(np.mean(diff.astype(np.float64) ** 2))
    v_out = (2.0 / diff.size) * diff
    v_xy, v_conic, v_col, v_op = rasterize_sum_backward(H, W, gids_s, bins, xys, conics, colors, opac, fidx, v_out)
    _, v_mean, v_scale, v_rot = project_gaussians_2d_scale_rot_backward(scales, rot, H, W, radii, conics, v_xy, v_conic)
    return loss, dict(v_means2d=v_mean, v_scale=v_scale, v_rot=v_rot, v_colors=v_col, v_opacity=v_op)


def scale_rot_radius_shadow(scales, rot, rel_tol=2e-5):
    """float64 SHADOW of the scale-rotation projection's radius (foward2d.cu:71-123 + helpers.cuh:45-68, the same
    formulas as gi2d_oracle_project_scale_rot_fwd, evaluated in double): returns (rf, near) where rf is the radius BEFORE
    the ceil and `near` flags the Gaussians whose rf lies within rel_tol of an integer.  sinf / cosf are libm here and
    libdevice on the GPU (1-2 ulp apart), so the fp32 radius may legitimately differ between the two ONLY where the ceil is
    that close to a step: tests compare the integer outputs exactly outside the flagged set."""
    s = np.asarray(scales, np.float64).reshape(-1, 2)
    th = np.asarray(rot, np.float64).reshape(-1)
    cs, sn = np.cos(th), np.sin(th)
    p, u, q, v = cs * s[:, 0], -(sn * s[:, 0]), sn * s[:, 1], cs * s[:, 1]
    a, b, c = p * p + q * q, p * u + q * v, u * u + v * v
    det = a * c - b * b
    bh = 0.5 * (a + c)
    disc = np.maximum(bh * bh - det, 0.1)
    rf = 3.0 * np.sqrt(bh + np.sqrt(disc))
    near = np.abs(rf - np.round(rf)) <= rel_tol * np.maximum(rf, 1.0)
    return rf, near
