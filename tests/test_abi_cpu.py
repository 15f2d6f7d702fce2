"""The C-ABI library loads without a GPU and exports exactly what include/gi2d.h declares (no compute calls here)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared():
    text = (ROOT / "include" / "gi2d.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gi2d_\w+)\s*\(", text)))


def test_header_declares_the_hot_path_entry_points():
    names = declared()
    for must in ("gi2d_project_cholesky_fwd", "gi2d_project_cholesky_bwd", "gi2d_project_scale_rot_fwd", "gi2d_project_scale_rot_bwd",
                 "gi2d_cov2d_bounds", "gi2d_scan_tiles", "gi2d_scan_tiles_static", "gi2d_emit_isects", "gi2d_sort_isects", "gi2d_tile_bins", "gi2d_bin_fused",
                 "gi2d_raster_sum_fwd", "gi2d_raster_sum_bwd", "gi2d_raster_sum_fwd_packed", "gi2d_raster_sum_bwd_packed"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from gaussianimage_b200 import _lib
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in declared() if not hasattr(lib, n)]
    assert not missing, f"declared in include/gi2d.h but not exported by libgi2d.so: {missing}"
    assert set(_lib.EXPORTS) == set(declared()), "gaussianimage_b200/_lib.py signatures out of sync with include/gi2d.h"
    assert _lib.lib().gi2d_abi_version() == 2
    assert _lib.lib().gi2d_status_string(0) == b"ok" and _lib.lib().gi2d_status_string(2) == b"workspace too small"


def test_workspace_queries_are_host_only_and_monotone():
    from gaussianimage_b200._lib import lib
    L = lib()
    a = L.gi2d_bin_fused_workspace_bytes(70000, 48, 32, 280000)
    b = L.gi2d_bin_fused_workspace_bytes(70000, 48, 32, 560000)
    assert 0 < a < b
    assert L.gi2d_sort_workspace_bytes(1000, 32, 43) < L.gi2d_sort_workspace_bytes(1000, 0, 64)
    assert L.gi2d_scan_workspace_bytes(70000) > 0


def test_library_is_sm100a_with_tma_and_no_fallback():
    import shutil
    import subprocess
    from gaussianimage_b200 import _lib
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not available")
    out = subprocess.run(["cuobjdump", "-lelf", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out
    sass = subprocess.run(["cuobjdump", "-sass", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass, "the rasterizers must stage their batches with bulk TMA copies"


def test_host_glue_loads_binds_the_same_library_and_checks_inputs():
    """csrc/torch_glue.cpp: host-only C++ over the same C ABI (no compute here: CPU tensors must be refused loudly)."""
    import torch
    from gaussianimage_b200 import _lib
    _lib.build_glue()
    T = _lib.glue()
    assert T is not None and T.abi_version() == 2
    for name in ("project_fwd", "project_bwd_packed", "fused_forward", "raster_bwd_packed", "project_autograd", "rasterize",
                 "check_overflow", "get_capacity", "set_capacity", "stats"):
        assert callable(getattr(T, name))
    mapped = {line.split()[-1] for line in open("/proc/self/maps") if "libgi2d" in line}
    assert mapped == {str(_lib.LIB_PATH)}, mapped          # one copy of the product library (it owns device-side state)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        T.project_fwd(0, 4, torch.zeros(4, 2), torch.zeros(4, 3), None, 32, 32, 2, 2)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        T.raster_bwd_packed(4, 32, 32, torch.zeros(4, 2, dtype=torch.int32), torch.zeros(48), torch.zeros(32, 32, 3))


def test_cpp_autograd_switch_routes_the_operators_and_still_refuses_cpu_tensors():
    """GI2D_CPP_AUTOGRAD / set_cpp_autograd: the operators as C++ autograd Functions (the default when the glue is built) or as
    Python ones.  No compute without a GPU: either route must be taken and must fail loudly on CPU tensors."""
    import torch
    import gaussianimage_b200 as G
    from gaussianimage_b200 import _lib
    _lib.build_glue()
    import os
    assert _lib._cpp_autograd == (os.environ.get("GI2D_CPP_AUTOGRAD", "1") != "0")
    prev = _lib._cpp_autograd
    _lib.set_cpp_autograd(False)
    assert _lib.cpp_autograd() is None
    with pytest.raises(RuntimeError, match="means2d must be a CUDA tensor"):
        G.project_gaussians_2d(torch.zeros(4, 2), torch.ones(4, 3), 32, 32, (2, 2, 1))
    _lib.set_cpp_autograd(True)
    try:
        assert _lib.cpp_autograd() is _lib.glue()
        with pytest.raises(RuntimeError, match="means2d must be a CUDA tensor"):
            G.project_gaussians_2d(torch.zeros(4, 2), torch.ones(4, 3), 32, 32, (2, 2, 1))
        with pytest.raises(RuntimeError, match="means2d must be a CUDA tensor"):
            G.project_gaussians_2d_scale_rot(torch.zeros(4, 2), torch.ones(4, 2), torch.ones(4, 1), 32, 32, (2, 2, 1))
        # CPU tensors never reach the fused chain: rasterize() declines (not the output of one of our projections), the staged
        # sequence then refuses the CPU tensor
        z = torch.zeros(4, 2)
        assert _lib.glue().rasterize(z, torch.zeros(4), torch.zeros(4, dtype=torch.int32), torch.zeros(4, 3), torch.zeros(4, dtype=torch.int32),
                                     torch.zeros(4, 3), torch.ones(4, 1), torch.ones(3), 32, 32, False) == []
        with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
            G.rasterize_gaussians_sum(z, torch.zeros(4), torch.zeros(4, dtype=torch.int32), torch.zeros(4, 3), torch.zeros(4, dtype=torch.int32),
                                      torch.zeros(4, 3), torch.ones(4, 1), 32, 32)
    finally:
        _lib.set_cpp_autograd(prev)
