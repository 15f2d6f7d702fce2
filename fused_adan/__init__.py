"""Top-level `fused_adan`: the module name the reference's optimizer imports for Adan(fused=True)
(optimizer.py:385, :427, :453).  Re-exports gaussianimage_b200.fused_adan (one CUDA kernel over the C ABI)."""
from gaussianimage_b200.fused_adan import adan_multi_tensor, adan_single_tensor  # noqa: F401

__all__ = ["adan_multi_tensor", "adan_single_tensor"]
