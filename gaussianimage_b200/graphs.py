"""CUDA-graph helper for callers that replay one whole iteration (fit.GraphFitter, bench.py): capture as torch does, then
post-process the captured graph before it is instantiated.

The only post-processing: small MEMSET nodes become fill-kernel nodes (include/gi2d.h: gi2d_graph_fold_memsets).  torch's
reduction kernels (`F.mse_loss`'s mean) clear a semaphore buffer with cudaMemsetAsync in front of the kernel; in a replayed
graph a memset node between kernel nodes costs several microseconds of idle time on either side (CUPTI timeline of the C2 fit
step: 2.4 us before, 4.4 us after a 1.4 us memset) -- a fill kernel of the same bytes does not.  The graph's work is unchanged."""
from __future__ import annotations

import ctypes as C
import os

import torch

from ._lib import check, lib


def new_graph() -> torch.cuda.CUDAGraph:
    """A CUDAGraph that keeps its cudaGraph_t after capture, so that fold_memsets() can edit it before instantiation."""
    return torch.cuda.CUDAGraph(keep_graph=os.environ.get("GI2D_FOLD_MEMSETS", "1") != "0")


def fold_memsets(g: torch.cuda.CUDAGraph, max_bytes: int = 1 << 16) -> int:
    """Call after capture_end (leaving the `with torch.cuda.graph(g)` block) and before the first replay.  Returns the number of
    memset nodes replaced (0 when the graph was created without keep_graph)."""
    try:
        raw = g.raw_cuda_graph()
    except RuntimeError:
        return 0                    # keep_graph=False: already instantiated by capture_end
    n = C.c_int(0)
    check(lib().gi2d_graph_fold_memsets(int(raw), max_bytes, C.addressof(n)), "graph_fold_memsets")
    g.instantiate()
    return int(n.value)
