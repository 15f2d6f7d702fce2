"""ctypes front of oracle/_ref/libgsplat_ref_{jit,fast}.so: the REFERENCE's own CUDA kernels (compiled by
oracle/Makefile from /root/reference, never copied) behind raw-pointer launchers (oracle/ref_harness.cu).
Test infrastructure only.  Kernels run on the legacy default stream; every call here synchronises."""
import ctypes as C
from pathlib import Path

import torch

_DIR = Path(__file__).resolve().parent.parent / "oracle" / "_ref"
_libs = {}


def available(flavor="jit"):
    return (_DIR / f"libgsplat_ref_{flavor}.so").exists()


def lib(flavor="jit"):
    if flavor not in _libs:
        _libs[flavor] = C.CDLL(str(_DIR / f"libgsplat_ref_{flavor}.so"))
    return _libs[flavor]


def _p(t):
    return C.c_void_p(t.data_ptr())


def _ok(rc, what):
    torch.cuda.synchronize()
    assert rc == 0, f"reference kernel {what} failed with cudaError {rc}"


def project_cholesky_fwd(means2d, L, H, W, flavor="jit"):
    N, dev = means2d.shape[0], means2d.device
    TX, TY = (W + 15) // 16, (H + 15) // 16
    xys, depths = torch.zeros(N, 2, device=dev), torch.zeros(N, device=dev)
    radii = torch.zeros(N, dtype=torch.int32, device=dev)
    conics, nth = torch.zeros(N, 3, device=dev), torch.zeros(N, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_project_cholesky_fwd(N, _p(means2d), _p(L), H, W, TX, TY, C.c_float(0.01), _p(xys), _p(depths),
                                             _p(radii), _p(conics), _p(nth)), "project_cholesky_fwd")
    return xys, depths, radii, conics, nth


def project_scale_rot_fwd(means2d, scales, rot, H, W, flavor="jit"):
    N, dev = means2d.shape[0], means2d.device
    TX, TY = (W + 15) // 16, (H + 15) // 16
    xys, depths = torch.zeros(N, 2, device=dev), torch.zeros(N, device=dev)
    radii = torch.zeros(N, dtype=torch.int32, device=dev)
    conics, nth = torch.zeros(N, 3, device=dev), torch.zeros(N, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_project_scale_rot_fwd(N, _p(means2d), _p(scales), _p(rot), H, W, TX, TY, C.c_float(0.01), _p(xys),
                                              _p(depths), _p(radii), _p(conics), _p(nth)), "project_scale_rot_fwd")
    return xys, depths, radii, conics, nth


def project_cholesky_bwd(means2d, L, H, W, radii, conics, v_xy, v_conic, flavor="jit"):
    N, dev = L.shape[0], L.device
    v_cov, v_mean, v_L = torch.zeros(N, 3, device=dev), torch.zeros(N, 2, device=dev), torch.zeros(N, 3, device=dev)
    v_depth = torch.zeros(N, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_project_cholesky_bwd(N, _p(means2d), _p(L), H, W, _p(radii), _p(conics), _p(v_xy), _p(v_depth),
                                             _p(v_conic), _p(v_cov), _p(v_mean), _p(v_L)), "project_cholesky_bwd")
    return v_cov, v_mean, v_L


def project_scale_rot_bwd(means2d, scales, rot, H, W, radii, conics, v_xy, v_conic, flavor="jit"):
    N, dev = scales.shape[0], scales.device
    v_cov, v_mean = torch.zeros(N, 3, device=dev), torch.zeros(N, 2, device=dev)
    v_scale, v_rot = torch.zeros(N, 2, device=dev), torch.zeros(N, 1, device=dev)
    v_depth = torch.zeros(N, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_project_scale_rot_bwd(N, _p(means2d), _p(scales), _p(rot), H, W, _p(radii), _p(conics), _p(v_xy),
                                              _p(v_depth), _p(v_conic), _p(v_cov), _p(v_mean), _p(v_scale), _p(v_rot)),
        "project_scale_rot_bwd")
    return v_cov, v_mean, v_scale, v_rot


def bin_and_sort(xys, depths, radii, nth, H, W, flavor="jit"):
    """The reference's Python staging (utils.py:116-167) around ITS kernels: cumsum, map, torch.sort, gather, bins."""
    N, dev = xys.shape[0], xys.device
    TX, TY = (W + 15) // 16, (H + 15) // 16
    cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
    n = int(cum[-1].item())
    ids = torch.zeros(n, dtype=torch.int64, device=dev)
    gids = torch.zeros(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_map_gaussian_to_intersects(N, _p(xys), _p(depths), _p(radii), _p(cum), TX, TY, _p(ids), _p(gids)),
        "map_gaussian_to_intersects")
    ids_s, order = torch.sort(ids)
    gids_s = torch.gather(gids, 0, order)
    bins = torch.zeros(max(n, TX * TY), 2, dtype=torch.int32, device=dev)   # reference: (n,2) -- Q4; enlarged to be safe
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_get_tile_bin_edges(n, _p(ids_s), _p(bins)), "get_tile_bin_edges")
    return n, cum, ids, gids, ids_s, gids_s.contiguous(), bins


def rasterize_sum_fwd(H, W, gids_s, bins, xys, conics, colors, opac, flavor="jit"):
    dev = xys.device
    TX, TY = (W + 15) // 16, (H + 15) // 16
    out, Ts = torch.zeros(H, W, 3, device=dev), torch.zeros(H, W, device=dev)
    fidx = torch.zeros(H, W, dtype=torch.int32, device=dev)
    bg = torch.ones(3)                      # host: the harness passes it by value
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_rasterize_sum_fwd(TX, TY, W, H, _p(gids_s), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac), _p(bg),
                                          _p(out), _p(Ts), _p(fidx)), "rasterize_sum_fwd")
    return out, Ts, fidx


def rasterize_sum_bwd(H, W, gids_s, bins, xys, conics, colors, opac, Ts, fidx, v_out, flavor="jit"):
    N, dev = xys.shape[0], xys.device
    TX, TY = (W + 15) // 16, (H + 15) // 16
    v_xy, v_conic = torch.zeros(N, 2, device=dev), torch.zeros(N, 3, device=dev)
    v_col, v_op = torch.zeros(N, 3, device=dev), torch.zeros(N, 1, device=dev)
    bg, v_alpha = torch.ones(3), torch.zeros(H, W, device=dev)
    torch.cuda.synchronize()
    _ok(lib(flavor).ref_rasterize_sum_bwd(TX, TY, W, H, _p(gids_s), _p(bins), _p(xys), _p(conics), _p(colors), _p(opac), _p(bg),
                                          _p(Ts), _p(fidx), _p(v_out), _p(v_alpha), _p(v_xy), _p(v_conic), _p(v_col), _p(v_op)),
        "rasterize_sum_bwd")
    return v_xy, v_conic, v_col, v_op
