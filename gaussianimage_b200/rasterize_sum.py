"""rasterize_gaussians_sum: tile binning + accumulated-sum rasterization, forward and backward.

Mirror of the reference operator gsplat/gsplat/rasterize_sum.py:14-253 (same name, arguments, defaults, return
value, gradient contract and Python-level errors).  Two execution paths with identical results:

* fast path (depths produced by one of this package's 2D projections, i.e. known to be all zero): the fused,
  sync-free chain  gi2d_bin_fused (one cooperative counting-sort kernel up to 4096 tiles; count / scan / scatter /
  per-tile sort above; scan + onesweep passes as the general fallback -- all ending in tile_bins + packed records)
  -> TMA-staged rasterizer.  The intersection count never leaves the device inside the chain; in the default *strict* mode it is
  read back once, AFTER everything has been enqueued (so the GPU is never idle waiting for the host, unlike the
  reference's mid-pipeline `.item()`, quirk Q12), to detect workspace overflow and retry.  While a CUDA graph is
  being captured, or after `set_sync_free(True)`, no read-back happens at all.
* general path (arbitrary depths): the reference's own staging -- cumulative intersects (one host read), 64-bit
  key emit, full 64-bit stable sort, bin edges, index-indirect rasterizer.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
from torch import Tensor
from torch.autograd import Function

from . import cuda as _C
import ctypes as C

from ._lib import BinArgs, check, cpp_autograd, glue, lib
from .utils import _depths_known_zero, bin_and_sort_gaussians, compute_cumulative_intersects

_sync_free = os.environ.get("GI2D_SYNC_FREE", "0") == "1"
_capacity: dict = {}        # (device index, N, TX, TY) -> intersection capacity of the fused workspace (Python Functions)
_last_status: dict = {}     # device index -> (scratch, status offset, capacity, key) of the most recent fused call
_sticky: dict = {}          # (device index, N, TX, TY) -> int32[2] device {max n_isect seen, overflowed}: only the host clears it


def set_sync_free(flag: bool) -> None:
    """True: never read the intersection count back (graph-capturable, maximum throughput); the caller is expected
    to call `check_overflow()` now and then.  False (default): strict mode -- the workspace is sized from the intersection
    total (C++ operators: published by the projection kernel; Python Functions: read back once at the end of forward)."""
    global _sync_free
    _sync_free = bool(flag)
    T = glue()
    if T is not None:
        T.set_sync_free(_sync_free)


def get_capacity(device_index: int, N: int, TX: int, TY: int) -> int:
    """Learned intersection capacity of this shape (0 = none yet) -- what sync-free / captured calls allocate."""
    T = cpp_autograd()
    c = T.get_capacity(device_index, N, TX, TY) if T is not None else 0
    return max(c, _capacity.get((device_index, N, TX, TY), 0))


def set_capacity(device_index: int, N: int, TX: int, TY: int, cap: int) -> None:
    _capacity[(device_index, N, TX, TY)] = int(cap)
    T = glue()
    if T is not None:
        T.set_capacity(device_index, N, TX, TY, int(cap))


def _overflowed(n: int, cap: int):
    raise RuntimeError(f"gi2d: {n} tile intersections exceeded the workspace capacity {cap}; "
                       "the capacity has been raised, re-run the step")


def check_overflow(device=None) -> int:
    """Synchronising check of `device`: the most recent fused forward AND the sticky status of every shape that ran
    sync-free or inside a replayed CUDA graph since the last check (a replay overwrites the per-call status; the sticky
    pair is only ever cleared here).  Returns the intersection count of the most recent call; raises if anything did not
    fit its workspace (those results were truncated) -- after raising the capacity of that shape and dropping the memoised
    binning, so that simply re-running (re-capturing) the step is correct."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    n_last = 0
    T = glue()
    if T is not None:
        n_last, over, n_over, cap = T.check_overflow(idx)
        if over:
            _geometry.clear()
            _overflowed(n_over, cap)
    for key, st in list(_sticky.items()):
        if key[0] != idx:
            continue
        n, over = st.tolist()
        if over:
            cap = _capacity.get(key, 0)
            _capacity[key] = max(cap, int(n * 1.25) + 1024)
            st.zero_()
            _geometry.clear()
            _last_status.pop(idx, None)
            _overflowed(n, cap)
    if idx in _last_status:
        scratch, o_status, cap, key = _last_status[idx]
        n, over = scratch[o_status:o_status + 8].view(torch.int32).tolist()
        if over:
            _capacity[key] = max(_capacity.get(key, 0), int(n * 1.25) + 1024)
            _geometry.clear()
            _last_status.pop(idx, None)
            _overflowed(n, cap)
        n_last = int(n) if not n_last else n_last
    return int(n_last)


def rasterize_gaussians_sum(
    xys: Tensor,
    depths: Tensor,
    radii: Tensor,
    conics: Tensor,
    num_tiles_hit: Tensor,
    colors: Tensor,
    opacity: Tensor,
    img_height: int,
    img_width: int,
    BLOCK_H: int = 16,
    BLOCK_W: int = 16,
    background: Optional[Tensor] = None,
    return_alpha: Optional[bool] = False,
) -> Tensor:
    """Returns out_img (H,W,3) -- or (out_img, out_alpha (H,W)) with return_alpha -- differentiable w.r.t. xys,
    conics, colors and opacity.  out_img is the plain accumulated sum (no transmittance, no background, Q6);
    with zero intersections it is `ones * background` as in the reference (rasterize_sum.py:121-129)."""
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255           # reference :53-55
    if background is not None:
        assert background.shape[0] == colors.shape[-1], (
            f"incorrect shape of background color tensor, expected shape {colors.shape[-1]}"
        )
    else:
        background = _default_background(colors)
    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise ValueError("xys must have dimensions (N, 2)")
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    T = cpp_autograd()
    if (T is not None and (BLOCK_H, BLOCK_W) == (16, 16) and colors.shape[-1] == 3 and xys.is_cuda
            and xys.dtype == torch.float32):
        # the whole fused path as ONE call into the host glue (csrc/torch_glue.cpp: rasterize); an empty result = these are
        # not the unmodified outputs of one of our projections -> the reference's staged sequence below
        if background.device != xys.device:
            background = background.to(xys.device)
        r = T.rasterize(xys.contiguous(), depths.contiguous(), radii.contiguous(), conics.contiguous(), num_tiles_hit.contiguous(),
                        colors.contiguous(), opacity.contiguous(), background.contiguous(), int(img_height), int(img_width),
                        bool(return_alpha))
        if r:
            return (r[0], r[1]) if return_alpha else r[0]
    return _RasterizeGaussiansSum.apply(
        xys.contiguous(), depths.contiguous(), radii.contiguous(), conics.contiguous(), num_tiles_hit.contiguous(),
        colors.contiguous(), opacity.contiguous(), img_height, img_width, BLOCK_H, BLOCK_W, background.contiguous(),
        return_alpha,
    )


_ones_cache: dict = {}


def _default_background(colors: Tensor) -> Tensor:
    """The reference allocates torch.ones(C) on every call (rasterize_sum.py:61-63); the value is a constant nobody
    writes to, so one tensor per (device, C) is kept (saves an allocation and a fill kernel per call)."""
    key = (colors.device, colors.shape[-1])
    bg = _ones_cache.get(key)
    if bg is None:
        bg = torch.ones(colors.shape[-1], dtype=torch.float32, device=colors.device)
        if not (colors.is_cuda and torch.cuda.is_current_stream_capturing()):     # a graph-pool tensor must not outlive its graph
            _ones_cache[key] = bg
    return bg


def _align(v: int, a: int = 256) -> int:
    return (v + a - 1) // a * a


class _RasterizeGaussiansSum(Function):
    @staticmethod
    def forward(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity, img_height, img_width,
                BLOCK_H=16, BLOCK_W=16, background=None, return_alpha=False):
        if (BLOCK_H, BLOCK_W) != (16, 16):
            raise ValueError("only BLOCK_H = BLOCK_W = 16 is coherent in the reference (helpers.cuh:36-41 divides by a "
                             "compile-time 16) and supported here")
        if colors.shape[-1] != 3:
            # the reference dispatches to _C.nd_rasterize_sum_forward, which is not bound (quirk Q10)
            raise AttributeError("nd_rasterize_sum_forward: only 3-channel colors are supported "
                                 "(the reference has no binding for other channel counts either)")
        H, W = int(img_height), int(img_width)
        TX, TY = (W + 15) // 16, (H + 15) // 16
        ctx.H, ctx.W, ctx.TX, ctx.TY = H, W, TX, TY
        ctx.set_materialize_grads(False)
        ctx.N = xys.size(0)
        fast = _depths_known_zero(depths) and xys.is_cuda and xys.dtype == torch.float32
        ctx.fast = fast
        global fast_calls, staged_calls
        if fast:
            fast_calls += 1
            out_img, final_Ts = _forward_fused(ctx, xys, radii, conics, num_tiles_hit, colors, opacity, background,
                                               return_alpha)
        else:
            staged_calls += 1
            out_img, final_Ts = _forward_staged(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity,
                                                background)
        if return_alpha:
            return out_img, 1 - final_Ts          # reference :181-183 (all zeros unless there were no intersections)
        return out_img

    @staticmethod
    def backward(ctx, v_out_img, v_out_alpha=None):
        N, H, W, TX, TY = ctx.N, ctx.H, ctx.W, ctx.TX, ctx.TY
        if v_out_img is None:                   # only out_alpha was used downstream: the reference ignores v_out_alpha (Q7)
            return (None,) * 13
        if ctx.fast:
            tile_bins, packed = ctx.saved_tensors
            dev = packed.device
            T = glue()
            if T is not None:
                grads = T.raster_bwd_packed(N, H, W, tile_bins, packed, v_out_img.contiguous())
            else:
                v_out = _C._f32(v_out_img.contiguous(), "v_out_img")
                with _C._guard(dev):
                    grads = torch.empty((N, 12), dtype=torch.float32, device=dev)     # 48-byte rows: vector reductions
                    check(lib().gi2d_raster_sum_bwd_packed(N, TX, TY, W, H, tile_bins.data_ptr(), packed.data_ptr(),
                                                           v_out.data_ptr(), grads.data_ptr(), _C._stream(dev)),
                          "rasterize_sum_backward")
            v_xy, v_conic, v_colors, v_opacity = grads[:, 0:2], grads[:, 2:5], grads[:, 5:8], grads[:, 8:9]
        else:
            gids_sorted, tile_bins, xys, conics, colors, opacity, background, final_Ts, final_idx = ctx.saved_tensors
            if ctx.num_intersects < 1:          # reference :209-213
                v_xy, v_conic = torch.zeros_like(xys), torch.zeros_like(conics)
                v_colors, v_opacity = torch.zeros_like(colors), torch.zeros_like(opacity)
            else:
                if v_out_alpha is None:
                    v_out_alpha = torch.zeros_like(v_out_img[..., 0])
                v_xy, v_conic, v_colors, v_opacity = _C.rasterize_sum_backward(
                    H, W, 16, 16, gids_sorted, tile_bins, xys, conics, colors, opacity, background, final_Ts, final_idx,
                    v_out_img, v_out_alpha)
        v_opacity = v_opacity.reshape(ctx.opacity_shape)
        return (v_xy, None, None, v_conic, None, v_colors, v_opacity, None, None, None, None, None, None)


class _HostStatus:
    """Ring of {n_isect, overflow} words in mapped pinned host memory.  The first kernel of the fused chain stores the
    pair there as soon as it knows it (csrc/binning.cu: gi2d_bin_fused_rows_notify) and the host polls the word
    instead of synchronising the stream: the reference's `.item()` (gsplat/gsplat/utils.py:117, quirk Q12) without the
    GPU ever waiting for the host, and without the host waiting for more than the scan kernel."""
    SLOTS = 64

    def __init__(self):
        self.t = torch.empty((self.SLOTS, 2), dtype=torch.int32).pin_memory()
        self.np = self.t.numpy()
        self.base = self.t.data_ptr()
        self.i = 0

    def acquire(self) -> int:
        slot = self.i
        self.i = (slot + 1) % self.SLOTS
        self.np[slot, 0] = -1
        return slot

    def wait(self, slot: int):
        a, spins = self.np, 0
        while a[slot, 0] == -1:
            spins += 1
            if spins % (1 << 20) == 0:            # seconds without an answer: surface a CUDA fault if there is one
                torch.cuda.synchronize()
                if a[slot, 0] == -1:
                    raise RuntimeError("gi2d: the binning chain never published its intersection count")
        return int(a[slot, 0]), int(a[slot, 1])


_host_status = None


class FusedResult:
    """Device tensors of one pass of the fused chain; gids_sorted / status are views made on demand."""
    __slots__ = ("out_img", "final_Ts", "final_idx", "tile_bins", "packed", "_scratch", "_o_gids", "_o_status", "_cap", "extra")

    @property
    def gids_sorted(self):
        return self._scratch[self._o_gids:self._o_gids + 4 * self._cap].view(torch.int32)

    @property
    def status(self):
        return self._scratch[self._o_status:self._o_status + 8].view(torch.int32)

    def __getitem__(self, k):           # dict-style access used by bench.py / tests
        return getattr(self, k)


def fused_chain(xys, radii, conics, num_tiles_hit, colors, opacity, background, H, W, cap, want_aux=False, host_slot=-1,
                raster=None, sticky=None, zero_g12=None):
    """One asynchronous pass of the fused chain (csrc/binning.cu gi2d_bin_fused + csrc/raster.cu packed forward) on the
    current stream with an intersection capacity `cap`.  Returns a FusedResult: out_img, final_Ts/final_idx (or None),
    tile_bins (T,2), packed (cap*12), gids_sorted (cap), status (2,) = [n, overflow].  host_slot >= 0 also publishes the
    status to that slot of the pinned host ring."""
    T = glue() if raster is None else None
    if T is not None:                      # the same allocations and two launches as one C++ call (csrc/torch_glue.cpp)
        r = FusedResult()
        (r.out_img, r.final_Ts, r.final_idx, r.tile_bins, r.packed, r._scratch, r._o_gids, r._o_status) = T.fused_forward(
            xys, radii, conics, num_tiles_hit, colors, opacity, background, H, W, cap,
            _host_status.base + 8 * host_slot if host_slot >= 0 else 0, bool(want_aux), 0 if sticky is None else sticky.data_ptr())
        r._cap = cap
        return r
    dev = xys.device
    L = lib()
    N = xys.size(0)
    TX, TY = (W + 15) // 16, (H + 15) // 16
    stream = _C._stream(dev)
    ws_bytes = L.gi2d_bin_fused_workspace_bytes(N, TX, TY, cap)
    # one scratch allocation: [workspace | gids_sorted | status]; tile_bins and packed are kept for backward
    o_gids = _align(ws_bytes)
    o_status = o_gids + _align(4 * cap)
    r = FusedResult()
    r._scratch = scratch = torch.empty((o_status + 256,), dtype=torch.uint8, device=dev)
    r._o_gids, r._o_status, r._cap = o_gids, o_status, cap
    r.tile_bins = tile_bins = torch.empty((TX * TY, 2), dtype=torch.int32, device=dev)
    r.packed = packed = torch.empty((cap * 12,), dtype=torch.float32, device=dev)
    base = scratch.data_ptr()
    host_ptr = _host_status.base + 8 * host_slot if host_slot >= 0 else 0
    a = BinArgs(N, TX, TY, 0, TY, cap, xys.data_ptr(), radii.data_ptr(), num_tiles_hit.data_ptr(), conics.data_ptr(),
                colors.data_ptr(), opacity.data_ptr(), base, ws_bytes, base + o_gids, tile_bins.data_ptr(), packed.data_ptr(),
                base + o_status, host_ptr or None, None if sticky is None else sticky.data_ptr(),
                None if zero_g12 is None else zero_g12.data_ptr())      # zero_g12: (N,12) gradient rows the chain zero-fills
    check(L.gi2d_bin_fused_ex(C.byref(a), stream), "bin_fused")
    if raster is not None:                 # another consumer of the lists (loss.py: forward that ends in the loss)
        r.out_img = r.final_Ts = r.final_idx = None
        raster(r, base + o_status, stream)
        return r
    r.out_img = out_img = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
    r.final_Ts = final_Ts = torch.empty((H, W), dtype=torch.float32, device=dev) if want_aux else None
    r.final_idx = final_idx = torch.empty((H, W), dtype=torch.int32, device=dev) if want_aux else None
    check(L.gi2d_raster_sum_fwd_packed(TX, TY, W, H, tile_bins.data_ptr(), packed.data_ptr(), base + o_status,
                                       background.data_ptr(), out_img.data_ptr(), _C._p(final_Ts),
                                       _C._p(final_idx), stream), "rasterize_sum_forward")
    return r


class _Geometry:
    """The last binning result per (device, stream): sorted ids, tile ranges, status.  rasterize_gaussians_sum called
    again with the SAME xys / radii / num_tiles_hit tensors (same storage, same version counters -- the reference model
    does it three times per forward, gaussianimage_cholesky.py:81,89,95, quirk Q13) only rebuilds the packed records from
    the new conics / colours / opacities (gi2d_repack_records) and rasterizes; the lists are bit-identical by construction."""
    __slots__ = ("xys", "radii", "nth", "versions", "shape", "result", "cap", "n_isect", "gen", "captured")

    def matches(self, xys, radii, nth, shape):
        # only tensors TAGGED by one of this package's projections (a generation number set where they were produced) can match:
        # a buffer somebody re-filled through a raw pointer has no tag of the same generation, whatever its address and version
        return (self.gen is not None and getattr(xys, "_gi2d_gen", None) == self.gen and self.shape == shape
                and self.xys.data_ptr() == xys.data_ptr() and self.radii.data_ptr() == radii.data_ptr()
                and self.nth.data_ptr() == nth.data_ptr() and self.versions == (xys._version, radii._version, nth._version))


_geometry: dict = {}
_memoise = True
geometry_hits = 0          # how many calls re-used a binning result (tests / bench read it)
fast_calls = 0             # calls that took the fused chain (depths known to be zero: produced by one of our projections)
staged_calls = 0           # calls that took the reference's staged sequence (arbitrary depths)


def stats() -> dict:
    """Counters of the operator's dispatch (tests read them): which path the calls took, how often the binning was re-used."""
    d = {"fast_calls": fast_calls, "staged_calls": staged_calls, "geometry_hits": geometry_hits}
    T = glue()
    if T is not None:
        for k, v in T.stats().items():
            d[k] = d.get(k, 0) + v
    return d


def set_memoise_binning(flag: bool):
    """Turn the re-use of the previous binning result for identical geometry tensors on / off (default on)."""
    global _memoise
    _memoise = bool(flag)
    _geometry.clear()
    T = glue()
    if T is not None:
        T.set_memoise(_memoise)


def fused_reraster(g, xys, conics, colors, opacity, background, H, W, want_aux=False, raster=None, zero_g12=None):
    """Rasterize the geometry of a previous fused_chain result with other conics / colours / opacities."""
    dev = xys.device
    L = lib()
    TX, TY = (W + 15) // 16, (H + 15) // 16
    stream = _C._stream(dev)
    src, cap = g.result, g.cap
    r = FusedResult()
    r._scratch, r._o_gids, r._o_status, r._cap = src._scratch, src._o_gids, src._o_status, cap
    r.tile_bins = src.tile_bins
    r.packed = packed = torch.empty((cap * 12,), dtype=torch.float32, device=dev)
    base = src._scratch.data_ptr()
    check(L.gi2d_repack_records_ex(cap, base + src._o_gids, base + src._o_status, xys.data_ptr(), conics.data_ptr(),
                                   colors.data_ptr(), opacity.data_ptr(), packed.data_ptr(), xys.size(0),
                                   None if zero_g12 is None else zero_g12.data_ptr(), stream), "repack_records")
    if raster is not None:
        r.out_img = r.final_Ts = r.final_idx = None
        raster(r, base + src._o_status, stream)
        return r
    r.out_img = out_img = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
    r.final_Ts = final_Ts = torch.empty((H, W), dtype=torch.float32, device=dev) if want_aux else None
    r.final_idx = final_idx = torch.empty((H, W), dtype=torch.int32, device=dev) if want_aux else None
    check(L.gi2d_raster_sum_fwd_packed(TX, TY, W, H, r.tile_bins.data_ptr(), packed.data_ptr(), base + src._o_status,
                                       background.data_ptr(), out_img.data_ptr(), _C._p(final_Ts),
                                       _C._p(final_idx), stream), "rasterize_sum_forward")
    return r


def _forward_fused(ctx, xys, radii, conics, num_tiles_hit, colors, opacity, background, want_aux, raster=None, zero_g12=None):
    global _host_status, geometry_hits
    N, H, W, TX, TY = ctx.N, ctx.H, ctx.W, ctx.TX, ctx.TY
    dev = xys.device
    xys, conics, colors = _C._f32(xys, "xys"), _C._f32(conics, "conics"), _C._f32(colors, "colors")
    opacity, background = _C._f32(opacity, "opacity"), _C._f32(background.to(dev), "background")
    radii, nth = _C._i32(radii, "radii"), _C._i32(num_tiles_hit, "num_tiles_hit")
    ctx.opacity_shape = opacity.shape
    key = (dev.index, N, TX, TY)
    strict = not (_sync_free or torch.cuda.is_current_stream_capturing())
    if strict and _host_status is None:
        _host_status = _HostStatus()
    capturing = torch.cuda.is_current_stream_capturing()
    sticky = _sticky.get(key)
    if sticky is None and not capturing:      # allocated (zeroed) at the first eager call: a capture only ever records its address
        sticky = _sticky[key] = torch.zeros(2, dtype=torch.int32, device=dev)
    with _C._guard(dev):
        gkey = (dev.index, _C._stream(dev))
        g = _geometry.get(gkey) if _memoise else None
        if g is not None and g.captured and not capturing:      # recorded inside a capture: its tensors live in that graph's pool
            _geometry.pop(gkey, None)
            g = None
        if g is not None and g.matches(xys, radii, nth, key):
            geometry_hits += 1
            r = fused_reraster(g, xys, conics, colors, opacity, background, H, W, want_aux, raster, zero_g12)
            if g.n_isect is not None:
                ctx.num_intersects = g.n_isect
            if raster is not None:
                return r
            ctx.save_for_backward(r.tile_bins, r.packed)
            return r.out_img, r.final_Ts
        while True:
            cap = _capacity.get(key)
            if cap is None:
                cap = _capacity[key] = max(4 * N, 1 << 14)
            slot = _host_status.acquire() if strict else -1
            r = fused_chain(xys, radii, conics, nth, colors, opacity, background, H, W, cap, want_aux, slot, raster,
                            None if strict else sticky, zero_g12)
            _last_status[dev.index] = (r._scratch, r._o_status, cap, key)
            n = None
            if not strict:
                break
            n, over = _host_status.wait(slot)      # published by the scan kernel; everything is already enqueued
            if not over:
                ctx.num_intersects = n
                break
            _capacity[key] = int(n * 1.25) + 1024  # truncated: grow and redo
        if _memoise and getattr(xys, "_gi2d_gen", None) is not None:
            g = _Geometry()
            g.xys, g.radii, g.nth, g.shape = xys.detach(), radii.detach(), nth.detach(), key   # no autograd graph kept alive
            g.versions = (xys._version, radii._version, nth._version)
            g.gen, g.captured = xys._gi2d_gen, capturing
            keep = FusedResult()               # lists only: the image is an autograd output and would pin the whole graph
            keep._scratch, keep._o_gids, keep._o_status, keep._cap, keep.tile_bins = r._scratch, r._o_gids, r._o_status, cap, r.tile_bins
            g.result, g.cap, g.n_isect = keep, cap, n
            _geometry[gkey] = g
    if raster is not None:
        return r
    ctx.save_for_backward(r.tile_bins, r.packed)
    return r.out_img, r.final_Ts


def _forward_staged(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity, background):
    """The reference's own sequence (rasterize_sum.py:109-179) over the staged entry points."""
    N, H, W, TX, TY = ctx.N, ctx.H, ctx.W, ctx.TX, ctx.TY
    ctx.opacity_shape = opacity.shape
    tile_bounds = (TX, TY, 1)
    num_intersects, cum_tiles_hit = compute_cumulative_intersects(num_tiles_hit)
    dev = xys.device
    if num_intersects < 1:
        out_img = torch.ones(H, W, colors.shape[-1], device=dev) * background
        gids_sorted = torch.zeros(0, 1, device=dev)
        tile_bins = torch.zeros(0, 2, device=dev)
        final_Ts = torch.zeros(H, W, device=dev)
        final_idx = torch.zeros(H, W, device=dev)
    else:
        _, _, _, gids_sorted, tile_bins = bin_and_sort_gaussians(N, num_intersects, xys, depths, radii, cum_tiles_hit,
                                                                 tile_bounds)
        out_img, final_Ts, final_idx = _C.rasterize_sum_forward(tile_bounds, (16, 16, 1), (W, H, 1), gids_sorted, tile_bins,
                                                                xys, conics, colors, opacity, background)
    ctx.num_intersects = num_intersects
    ctx.save_for_backward(gids_sorted, tile_bins, xys, conics, colors, opacity, background, final_Ts, final_idx)
    return out_img, final_Ts
