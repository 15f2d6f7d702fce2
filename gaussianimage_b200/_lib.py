"""ctypes loader of libgi2d.so (the C ABI in include/gi2d.h).

There is NO fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libgi2d.so"
_lib = None

_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t

# name -> (restype, argtypes); pointers are passed as integers (tensor.data_ptr()), streams as integers too.
_SIGNATURES = {
    "gi2d_abi_version": (_i, []),
    "gi2d_status_string": (C.c_char_p, [_i]),
    "gi2d_project_cholesky_fwd": (_i, [_i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_decode_project_cholesky_fwd": (_i, [_i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_fwd": (_i, [_i, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_cholesky_bwd": (_i, [_i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_bwd": (_i, [_i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_cholesky_bwd_packed": (_i, [_i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_bwd_packed": (_i, [_i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_cov2d_bounds": (_i, [_i, _vp, _vp, _vp, _vp]),
    "gi2d_scan_workspace_bytes": (_sz, [_i]),
    "gi2d_scan_tiles": (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gi2d_emit_isects": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "gi2d_sort_workspace_bytes": (_sz, [_i, _i, _i]),
    "gi2d_sort_isects": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _vp, _sz, _vp]),
    "gi2d_tile_bins": (_i, [_i, _vp, _i, _vp, _vp]),
    "gi2d_bin_fused_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "gi2d_bin_fused_kernel_count": (_i, [_i, _i, _i]),
    "gi2d_adan_multi_tensor": (_i, [_i] + [_vp] * 7 + [C.c_double] * 9 + [_i, C.c_double, _i, _vp]),
    "gi2d_adan_state_bytes": (_sz, []),
    "gi2d_adan_state_init": (_i, [_vp] + [C.c_double] * 7 + [_i]),
    "gi2d_adan_multi_tensor_graph": (_i, [_i] + [_vp] * 7 + [_vp, _i, _vp]),
    "gi2d_raster_mse_workspace_bytes": (_sz, []),
    "gi2d_raster_sum_fwd_packed_mse": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gi2d_raster_sum_bwd_packed_scaled": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_repack_records": (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused_rows": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused_rows_notify": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed_rows": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed_rows_mc": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_bwd": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_bwd_packed": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
}
EXPORTS = tuple(_SIGNATURES)


def build(verbose: bool = False) -> Path:
    """Compile libgi2d.so in-tree with nvcc for sm_100a (gaussianimage_b200/csrc/Makefile)."""
    r = subprocess.run(["make", "-C", str(_PKG / "csrc"), "-j4"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libgi2d.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} is missing: the gi2d CUDA library has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback."
            )
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)      # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        if L.gi2d_abi_version() != 1:
            raise RuntimeError("libgi2d.so ABI version mismatch")
        _lib = L
    return _lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().gi2d_status_string(status).decode()
        raise RuntimeError(f"{what} failed: {msg} (status {status})")
