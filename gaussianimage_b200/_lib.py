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
    "gi2d_graph_fold_memsets": (_i, [_vp, _sz, _vp]),
    "gi2d_project_cholesky_fwd": (_i, [_i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_decode_project_cholesky_fwd": (_i, [_i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_fwd": (_i, [_i, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_cholesky_bwd": (_i, [_i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_bwd": (_i, [_i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_cholesky_bwd_packed": (_i, [_i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_scale_rot_bwd_packed": (_i, [_i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_project_fwd_total": (_i, [_i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_cov2d_bounds": (_i, [_i, _vp, _vp, _vp, _vp]),
    "gi2d_scan_workspace_bytes": (_sz, [_i]),
    "gi2d_scan_tiles": (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gi2d_scan_tiles_static": (_i, [_i, _vp, _vp, _vp, _vp]),
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
    "gi2d_repack_records_ex": (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "gi2d_bin_fused_ex": (_i, [_vp, _vp]),
    "gi2d_raster_sum_bwd_packed_acc": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused_rows": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_bin_fused_rows_notify": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed_rows": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed_rows_mc": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed_rows_p2p": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "gi2d_raster_sum_fwd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_fwd_packed": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_bwd": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gi2d_raster_sum_bwd_packed": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
}
EXPORTS = tuple(_SIGNATURES)


class BinArgs(C.Structure):
    """struct gi2d_bin_args of include/gi2d.h."""
    _fields_ = [("N", _i), ("TX", _i), ("TY", _i), ("ty_begin", _i), ("ty_end", _i), ("capacity", _i),
                ("xys", _vp), ("radii", _vp), ("num_tiles_hit", _vp), ("conics", _vp), ("colors", _vp), ("opacities", _vp),
                ("ws", _vp), ("ws_bytes", _sz), ("gids_sorted", _vp), ("tile_bins", _vp), ("packed", _vp),
                ("status_out", _vp), ("status_host", _vp), ("status_sticky", _vp), ("zero_grads12", _vp)]


def build(verbose: bool = False) -> Path:
    """Compile libgi2d.so in-tree with nvcc for sm_100a (gaussianimage_b200/csrc/Makefile)."""
    r = subprocess.run(["make", "-C", str(_PKG / "csrc"), "-j4"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libgi2d.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


GLUE_PATH = _PKG / "_gi2d_torch.so"
_glue = False               # False = not looked up yet, None = absent / switched off


def build_glue() -> Path:
    """Compile csrc/torch_glue.cpp (host-only C++: allocation + launch sequences over libgi2d.so, no device code) against
    this interpreter's torch into gaussianimage_b200/_gi2d_torch.so (in-tree; links libgi2d.so through $ORIGIN)."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension as E
    src = _PKG / "csrc" / "torch_glue.cpp"
    hdr = _PKG.parent / "include" / "gi2d.h"
    if GLUE_PATH.exists() and GLUE_PATH.stat().st_mtime >= max(src.stat().st_mtime, hdr.stat().st_mtime, LIB_PATH.stat().st_mtime):
        return GLUE_PATH
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-w", str(src), "-o", str(GLUE_PATH),
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-DTORCH_EXTENSION_NAME=_gi2d_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-I{_PKG.parent / 'include'}", f"-I{sysconfig.get_paths()['include']}",
           f"-I{cuda_home}/include"]
    cmd += [f"-I{d}" for d in E.include_paths()]
    for d in E.library_paths():
        cmd += [f"-L{d}", f"-Wl,-rpath,{d}"]
    cmd += ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
            f"-L{_PKG}", "-lgi2d", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building _gi2d_torch.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return GLUE_PATH


def glue():
    """The optional C++ host glue module (csrc/torch_glue.cpp), or None: absent, or GI2D_HOST_GLUE=0.  Both host paths
    drive the same kernels through the same C ABI; this one only shortens the host side of an eager call."""
    global _glue
    if _glue is False:
        _glue = None
        if os.environ.get("GI2D_HOST_GLUE", "1") != "0" and GLUE_PATH.exists():
            lib()                                   # libgi2d.so first, by path: the glue's DT_NEEDED then binds to this copy
            import importlib.util

            import torch  # noqa: F401  (libtorch must be loaded before the extension)
            try:
                spec = importlib.util.spec_from_file_location("gaussianimage_b200._gi2d_torch", str(GLUE_PATH))
                mod = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(mod)
            except (ImportError, OSError) as e:     # e.g. built against another torch: the ctypes host path runs instead
                import warnings
                warnings.warn(f"gi2d: host glue {GLUE_PATH.name} could not be loaded ({e}); using the ctypes host path "
                              "(same kernels). Rebuild with `python -c 'import __graft_entry__ as g; g.build()'`.")
                return None
            if mod.abi_version() != 2:
                raise RuntimeError("_gi2d_torch.so was built against another libgi2d ABI")
            _glue = mod
    return _glue


_cpp_autograd = os.environ.get("GI2D_CPP_AUTOGRAD", "1") != "0"


def cpp_autograd():
    """The glue module when the operators run as C++ autograd Functions with the fused chain's policy in C++
    (csrc/torch_glue.cpp: ProjectNode, RasterNode, rasterize) -- the default whenever the glue is built -- else None:
    GI2D_CPP_AUTOGRAD=0 / set_cpp_autograd(False) select the Python autograd Functions over the same kernels."""
    return glue() if _cpp_autograd else None


def set_cpp_autograd(flag: bool) -> None:
    global _cpp_autograd
    _cpp_autograd = bool(flag)


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
        if L.gi2d_abi_version() != 2:
            raise RuntimeError("libgi2d.so ABI version mismatch")
        _lib = L
    return _lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().gi2d_status_string(status).decode()
        raise RuntimeError(f"{what} failed: {msg} (status {status})")
