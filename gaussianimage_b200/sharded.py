"""Multi-GPU forms of the path (new functionality; the reference is single-GPU, SURVEY.md sections 2.3 / 8e).

* Image-parallel fitting (BASELINE config 4): every image is an independent optimisation, so ranks simply take images
  `rank, rank + world, ...` -- `image_shard()`.  No collective touches the hot path.
* Tile-row-sharded render-only decode of very large frames (BASELINE config 5): every rank holds all Gaussians,
  projects them, bins and rasterizes ONLY the tile rows it owns (csrc: gi2d_bin_fused_rows /
  gi2d_raster_sum_fwd_packed_rows) straight into its band of a full-size image buffer, and the frame is completed on
  every rank either by ONE in-place NCCL all-gather over NVLink (`fused=False`) or -- the B200/NVSwitch-native form,
  `fused=True` -- INSIDE the rasterizer: the frame buffer lives in torch symmetric memory and every pixel store is an
  NVLS multicast store (`multimem.st`, csrc/raster.cu) that the switch replicates into every rank's buffer while the
  remaining tiles are still being computed; two symmetric-memory barriers per frame replace the collective
  -- `RowShardedRenderer`.  Bands are equal
  (ceil(TY / world) tile rows, the buffer is padded to world * band rows) so that the all-gather needs no staging
  and no reassembly.  Per-tile order and per-pixel arithmetic do not depend on which rank owns a tile, so the
  gathered frame is bit-identical to the single-GPU one.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

import ctypes as C

from . import cuda as _C
from ._lib import BinArgs, check, lib


def image_shard(num_images: int, rank: int, world: int) -> List[int]:
    """Images fitted by `rank` (round-robin, like train.py:366-389's sequential loop dealt out to ranks)."""
    return list(range(rank, num_images, world))


def row_partition(tile_rows: int, world: int) -> List[Tuple[int, int]]:
    """Equal bands of ceil(tile_rows / world) tile rows; trailing ranks may get a short or empty band."""
    band = (tile_rows + world - 1) // world
    return [(min(r * band, tile_rows), min((r + 1) * band, tile_rows)) for r in range(world)]


class RowShardedRenderer:
    """Render-only (decode) path sharded by tile rows over the ranks of `group` (None = single process).

    fused=True (needs NVLS multicast): TWO symmetric frame buffers used alternately.  Frame i is rasterized with multicast
    stores into buffer i & 1 of EVERY rank; one symmetric-memory barrier per frame, issued on a side stream, marks it
    complete.  The rasterizer of frame i only waits for the barrier of frame i-1 -- which also says that every rank has
    finished its stream work up to its own frame i-1, i.e. has consumed frame i-2, the previous content of the buffer being
    overwritten -- so the projection and binning of frame i overlap with the stores of frame i-1 still draining, and a
    consumer may read frame i while frame i+1 is produced.  Contract: a frame returned by render*() stays valid until the
    second render*() call after it; consume it on the stream it was rendered on (or synchronise with that stream)."""

    def __init__(self, H: int, W: int, device, group: Optional[dist.ProcessGroup] = None, emulate_world: int = 0,
                 fused: bool = False, exchange: str = "multicast", ce_queues: int = 0):
        """exchange (fused mode): "multicast" = one NVLS `multimem.st` per 16 bytes from the rasterizer, replicated by the switch;
        "p2p" = plain stores from the rasterizer to the peer-mapped frame buffer of every other rank; "ce" = the rasterizer
        writes its band locally and the COPY ENGINES push it to every peer (peer-to-peer cudaMemcpyAsync on the side stream)
        while the SMs already work on the next frame."""
        self.H, self.W = int(H), int(W)
        self.TX, self.TY = (self.W + 15) // 16, (self.H + 15) // 16
        self.group = group
        self.world = dist.get_world_size(group) if (group is not None or dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        if emulate_world:           # tests: one process plays every rank in turn (set .rank, call render_band), no collective
            self.world, self.rank = emulate_world, 0
        self.bands = row_partition(self.TY, self.world)
        self.band_rows = (self.TY + self.world - 1) // self.world * 16          # pixel rows per (padded) band
        self.device = torch.device(device)
        # full frame padded to world equal bands; frame() is the (H,W,3) view the caller sees
        shape = (self.world * self.band_rows, self.W, 3)
        self.fused, self._mcs, self._hdls = False, [0], [None]
        if fused and self.world > 1 and not emulate_world:
            import torch.distributed._symmetric_memory as symm_mem
            self.buffers, self._hdls, self._mcs, self._peers = [], [], [], []
            self.exchange = exchange
            for _ in range(2):
                buf = symm_mem.empty(shape, dtype=torch.float32, device=self.device)
                hdl = symm_mem.rendezvous(buf, group if group is not None else dist.group.WORLD)
                mc = int(hdl.multicast_ptr or 0)
                if exchange == "multicast" and not mc:
                    raise RuntimeError("gi2d: fused row-sharded rendering needs NVLS multicast (symmetric memory multicast_ptr is null)")
                ptrs = [int(p) for k, p in enumerate(hdl.buffer_ptrs) if k != self.rank]
                if exchange == "p2p" and len(ptrs) > 7:
                    raise RuntimeError("gi2d: peer-store exchange supports at most 8 ranks")
                buf.zero_()
                self.buffers.append(buf); self._hdls.append(hdl); self._mcs.append(mc)
                self._peers.append((C.c_void_p * max(len(ptrs), 1))(*ptrs))
                if exchange == "ce":
                    self._peer_bufs = getattr(self, "_peer_bufs", [])
                    # rank r pushes to r+1, r+2, ...: at every step the copies of all ranks form a permutation, so no
                    # destination receives from two sources at once
                    self._peer_bufs.append([hdl.get_buffer((self.rank + k) % self.world, shape, torch.float32)
                                            for k in range(1, self.world)])
            self.fused = True
            self._side = torch.cuda.Stream(device=self.device)
            # Copy queues of the copy-engine exchange (the peers are dealt round-robin over them).  ONE by default: the copies of
            # a queue run one after the other, so at every moment the copies of all ranks form a permutation (rank r -> r + k) and no
            # destination takes two streams at once.  More queues break that: N=8 measured 555-563 us per frame with one queue,
            # 570-614 with three, 930-1030 with two, four or seven (tools/ce_queues.py); N=4: 436-441 with one or two.
            import os
            nq = int(ce_queues or os.environ.get("GI2D_CE_QUEUES", "1"))
            self._ce_extra = [torch.cuda.Stream(device=self.device) for _ in range(max(0, min(nq, self.world - 1) - 1))]
            self._done = [torch.cuda.Event(), torch.cuda.Event()]
            torch.cuda.synchronize(self.device)
            self._hdls[0].barrier(channel=0)            # every rank's buffers are zeroed before anybody's stores arrive
            for e in self._done:
                e.record(torch.cuda.current_stream(self.device))
        else:
            self.buffers = [torch.zeros(shape, dtype=torch.float32, device=self.device)]
        self._cur = 0               # buffer of the most recent frame
        self._frame_no = 0
        self.background = torch.ones(3, dtype=torch.float32, device=self.device)
        self._cap = None
        self.status = None
        self._sticky = None          # int32[2] {max intersections seen in this rank's band, overflowed}: cleared only by the host
        self._frames = 0
        self.check_every = 64        # frames between two reads of the sticky status once a capacity has proved sufficient

    @property
    def buffer(self) -> torch.Tensor:
        return self.buffers[self._cur]

    def frame(self) -> torch.Tensor:
        return self.buffer[: self.H]

    def render_band(self, xys, radii, conics, num_tiles_hit, colors, opacity) -> None:
        """Bin + rasterize this rank's tile rows into its band of the (next) frame buffer (asynchronous, current stream)."""
        r0, r1 = self.bands[self.rank]
        b = (self._cur ^ 1) if self.fused else 0
        self._cur = b
        self._frame_no += 1
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        if r0 >= r1:
            if self.fused:
                cur.wait_event(self._done[b] if self.exchange == "ce" else self._done[b ^ 1])
                self._post_barrier(b, cur)
            return
        L = lib()
        N = xys.size(0)
        if self._cap is None:
            self._cap = max(int(4 * N * (r1 - r0) / max(self.TY, 1)) + (1 << 14), 1 << 14)
        with _C._guard(dev):
            stream = _C._stream(dev)
            first = True
            while True:
                cap = self._cap
                ws_bytes = L.gi2d_bin_fused_workspace_bytes(N, self.TX, self.TY, cap)
                if getattr(self, "_ws_cap", None) != cap:       # persistent workspace (decode loops reuse it)
                    self._ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
                    self._gids = torch.empty((cap,), dtype=torch.int32, device=dev)
                    self._bins = torch.empty((self.TX * self.TY, 2), dtype=torch.int32, device=dev)
                    self._packed = torch.empty((cap * 12,), dtype=torch.float32, device=dev)
                    self.status = torch.zeros((2,), dtype=torch.int32, device=dev)
                    self._sticky = torch.zeros((2,), dtype=torch.int32, device=dev)
                    self._ws_cap = cap
                    self._cap_checked = False
                a = BinArgs(N, self.TX, self.TY, r0, r1, cap, xys.data_ptr(), radii.data_ptr(), num_tiles_hit.data_ptr(),
                            conics.data_ptr(), colors.data_ptr(), opacity.data_ptr(), self._ws.data_ptr(), ws_bytes,
                            self._gids.data_ptr(), self._bins.data_ptr(), self._packed.data_ptr(), self.status.data_ptr(), None,
                            self._sticky.data_ptr(), None)
                check(L.gi2d_bin_fused_ex(C.byref(a), stream), "bin_fused_rows")
                if self.fused and first:
                    # the frame buffer is touched from here on: the previous frame must be complete on every rank (and with it
                    # every rank's consumption of the frame before that, whose buffer this is).  Copy-engine exchange: only this
                    # rank's own band is written here, and only this rank's copies of frame i-2 still read it.
                    cur.wait_event(self._done[b] if self.exchange == "ce" else self._done[b ^ 1])
                    first = False
                # H here is the padded buffer height so that every pixel row of the band exists
                buf = self.buffers[b]
                if self.fused and self.exchange == "p2p":
                    check(L.gi2d_raster_sum_fwd_packed_rows_p2p(self.TX, self.TY, self.W, buf.shape[0], r0, r1, self._bins.data_ptr(),
                                                                self._packed.data_ptr(), buf.data_ptr(), self._peers[b], self.world - 1,
                                                                stream), "rasterize_sum_forward (rows, peer stores)")
                else:
                    mc = self._mcs[b] if (self.fused and self.exchange == "multicast") else 0
                    check(L.gi2d_raster_sum_fwd_packed_rows_mc(self.TX, self.TY, self.W, buf.shape[0], r0, r1, self._bins.data_ptr(),
                                                               self._packed.data_ptr(), 0, 0, buf.data_ptr(), mc, 0, 0, stream),
                          "rasterize_sum_forward (rows)")
                self._frames += 1
                if torch.cuda.is_current_stream_capturing():
                    break
                if getattr(self, "_cap_checked", False) and self._frames % self.check_every:
                    break
                # first frames of a capacity: this frame's own status; later, every `check_every` frames: the sticky pair, which
                # remembers an overflow of ANY frame since it was last cleared (a later frame or model may have more
                # intersections in this band than the one the capacity was learned on)
                n, over = (self._sticky if getattr(self, "_cap_checked", False) else self.status).tolist()
                if not over:
                    self._cap_checked = True
                    break
                self._sticky.zero_()
                self._cap = int(n * 1.25) + 1024
        if self.fused:
            self._post_barrier(b, cur)

    def _post_barrier(self, b: int, cur) -> None:
        """The one barrier of a fused frame, on the side stream: all ranks' multicast stores of this frame have landed."""
        ev = torch.cuda.Event()
        ev.record(cur)
        with torch.cuda.stream(self._side):
            self._side.wait_event(ev)
            if self.exchange == "ce":       # (the side stream is ordered: these copies follow the barrier of the previous frame)
                lo, hi = self.rank * self.band_rows, (self.rank + 1) * self.band_rows
                mine = self.buffers[b][lo:hi]
                peers = self._peer_bufs[b]
                nq = len(self._ce_extra) + 1
                if nq > 1:                  # several copy queues, so that several copy engines keep the links busy back to back
                    go = torch.cuda.Event()
                    go.record(self._side)
                    backs = []
                    for q, extra in enumerate(self._ce_extra, start=1):
                        with torch.cuda.stream(extra):
                            extra.wait_event(go)
                            for peer in peers[q::nq]:
                                peer[lo:hi].copy_(mine, non_blocking=True)
                            back = torch.cuda.Event()
                            back.record(extra)
                            backs.append(back)
                    for peer in peers[0::nq]:
                        peer[lo:hi].copy_(mine, non_blocking=True)
                    for back in backs:
                        self._side.wait_event(back)
                else:
                    for peer in peers:
                        peer[lo:hi].copy_(mine, non_blocking=True)
            self._hdls[b].barrier(channel=0)
            self._done[b].record(self._side)

    def gather(self) -> torch.Tensor:
        """Completes the frame on every rank: fused mode = the current stream waits for the frame's barrier (the multicast
        stores already delivered every band everywhere); otherwise an in-place all-gather (every rank contributes
        buffer[rank*band : (rank+1)*band])."""
        if self.fused:
            torch.cuda.current_stream(self.device).wait_event(self._done[self._cur])
        elif self.world > 1 and dist.is_initialized():
            mine = self.buffer[self.rank * self.band_rows: (self.rank + 1) * self.band_rows]
            dist.all_gather_into_tensor(self.buffer, mine, group=self.group)
        return self.frame()

    def render_compressed(self, xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks,
                          opacity) -> torch.Tensor:
        """decode step from the compressed representation (compress_wo_ec, gaussianimage_cholesky.py:154-159): the
        de-quantisation prologue and the projection are ONE kernel (decode.py), then own band, then the exchange."""
        from .decode import decode_project_gaussians_2d
        xys, depths, radii, conics, nth, colors = decode_project_gaussians_2d(
            xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks, self.H, self.W,
            (self.TX, self.TY, 1))
        self.render_band(xys, radii, conics, nth, colors, opacity)
        return self.gather()

    def render_compressed_async(self, xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks,
                                opacity) -> torch.Tensor:
        """render_compressed WITHOUT making the current stream wait for the frame's completion: for pipelines that keep several
        frames in flight (call gather() -- or wait_frames() at the end -- before reading the returned view)."""
        from .decode import decode_project_gaussians_2d
        xys, depths, radii, conics, nth, colors = decode_project_gaussians_2d(
            xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks, self.H, self.W,
            (self.TX, self.TY, 1))
        self.render_band(xys, radii, conics, nth, colors, opacity)
        if not self.fused:
            return self.gather()
        return self.frame()

    def wait_frames(self) -> None:
        """Current stream waits until every frame rendered so far is complete on every rank."""
        if self.fused:
            cur = torch.cuda.current_stream(self.device)
            for e in self._done:
                cur.wait_event(e)

    def render_compressed_band(self, xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks,
                               opacity) -> torch.Tensor:
        """Like render_compressed, but WITHOUT the exchange: returns only this rank's band of the frame (rows
        [rank*band_rows, ...)) -- for consumers that take the frame band by band (per-band encode, display walls)."""
        from .decode import decode_project_gaussians_2d
        xys, depths, radii, conics, nth, colors = decode_project_gaussians_2d(
            xyz, quant_cholesky_elements, scale, beta, cholesky_bound, feature_dc_index, codebooks, self.H, self.W,
            (self.TX, self.TY, 1))
        self.render_band(xys, radii, conics, nth, colors, opacity)
        return self.buffer[self.rank * self.band_rows: (self.rank + 1) * self.band_rows]

    def render(self, means2d, L_elements, colors, opacity) -> torch.Tensor:
        """decode step: project all Gaussians (every rank), render own band, all-gather -> (H,W,3) frame."""
        N = means2d.shape[0]
        xys, depths, radii, conics, nth = _C.project_gaussians_2d_forward(N, means2d, L_elements, self.H, self.W,
                                                                          (self.TX, self.TY, 1))
        self.render_band(xys, radii, conics, nth, colors, opacity)
        return self.gather()
