"""Driver for an `ncu --profile-from-start off` capture of exactly ONE launch of each kernel of the chain (binning, forward
and backward rasterizer, through the C ABI) at a given config, L2 flushed before each.
  python tools/prof_raster.py [C2]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200._lib import check, lib
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
s = make_config(cfg)
d = s.to("cuda:0")
dev = d.gt.device
H, W, (TX, TY, _) = s.H, s.W, s.tile_bounds
N = d.means2d.shape[0]
L = lib()
xys, depths, radii, conics, nth = G.project_gaussians_2d(d.means2d, d.cholesky, H, W, s.tile_bounds)
n = int(nth.sum().item())
cap = int(n * 1.25) + 1024
ws_bytes = L.gi2d_bin_fused_workspace_bytes(N, TX, TY, cap)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
gids = torch.empty(cap, dtype=torch.int32, device=dev)
bins = torch.empty(TX * TY, 2, dtype=torch.int32, device=dev)
packed = torch.empty(cap * 12, device=dev)
status = torch.empty(2, dtype=torch.int32, device=dev)
out = torch.empty(H, W, 3, device=dev)
bg = torch.ones(3, device=dev)
v_out = torch.randn(H, W, 3, device=dev) * 1e-3
grads = torch.empty(N, 12, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream


def binning():
    check(L.gi2d_bin_fused(N, TX, TY, xys.data_ptr(), radii.data_ptr(), nth.data_ptr(), conics.data_ptr(), d.colors.data_ptr(),
                           d.opacity.data_ptr(), cap, ws.data_ptr(), ws_bytes, gids.data_ptr(), bins.data_ptr(), packed.data_ptr(),
                           status.data_ptr(), st), "bin")


def fwd():
    check(L.gi2d_raster_sum_fwd_packed(TX, TY, W, H, bins.data_ptr(), packed.data_ptr(), status.data_ptr(), bg.data_ptr(),
                                       out.data_ptr(), 0, 0, st), "fwd")


def bwd():
    check(L.gi2d_raster_sum_bwd_packed(N, TX, TY, W, H, bins.data_ptr(), packed.data_ptr(), v_out.data_ptr(), grads.data_ptr(), st), "bwd")


for _ in range(3):
    binning(); fwd(); bwd()
torch.cuda.synchronize()
for fn in (binning, fwd, bwd):
    flush.zero_()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
print("ok", status.tolist())
