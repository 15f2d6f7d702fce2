"""Experiment: GPU time of the fused chain truncated after the scan kernel (GI2D_DBG bit 4), in a CUDA graph."""
import statistics, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import gaussianimage_b200 as G
from gaussianimage_b200._lib import check, lib
from gaussianimage_b200.synth import make_config
s = make_config("C2"); d = s.to("cuda:0"); dev = d.gt.device
H, W, (TX, TY, _) = s.H, s.W, s.tile_bounds
N = d.means2d.shape[0]; L = lib()
xys, depths, radii, conics, nth = G.project_gaussians_2d(d.means2d, d.cholesky, H, W, s.tile_bounds)
cap = 280000
ws_bytes = L.gi2d_bin_fused_workspace_bytes(N, TX, TY, cap)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev); gids = torch.empty(cap, dtype=torch.int32, device=dev)
bins = torch.empty(TX * TY, 2, dtype=torch.int32, device=dev); packed = torch.empty(cap * 12, device=dev)
status = torch.empty(2, dtype=torch.int32, device=dev)
def binning():
    check(L.gi2d_bin_fused(N, TX, TY, xys.data_ptr(), radii.data_ptr(), nth.data_ptr(), conics.data_ptr(), d.colors.data_ptr(),
                           d.opacity.data_ptr(), cap, ws.data_ptr(), ws_bytes, gids.data_ptr(), bins.data_ptr(), packed.data_ptr(),
                           status.data_ptr(), torch.cuda.current_stream().cuda_stream), "bin")
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    binning()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    binning()
ts = []
for _ in range(50):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
print(f"median {statistics.median(ts):.2f} us min {min(ts):.2f} (warm L2)")
