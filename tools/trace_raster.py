"""Per-warp timeline of the two rasterizers: uses a copy of the library built with -DGI2D_TRACE (gaussianimage_b200/_trace/,
never the product library), runs ONE launch of each at a config after the product's own projection, and prints where the
kernel's time goes: launch ramp (first entry -> a warp's entry), load latency (entry -> first chunk landed), evaluation, tail
(a warp's exit -> the last exit of the grid), per SM and over the grid.     python tools/trace_raster.py [C2] [--rebuild]"""
import ctypes as C
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import gaussianimage_b200 as G
from gaussianimage_b200.synth import make_config

out_dir = ROOT / "gaussianimage_b200" / "_trace"
out_dir.mkdir(exist_ok=True)
if not (out_dir / "libgi2d_trace.so").exists() or "--rebuild" in sys.argv:
    r = subprocess.run(["make", "-C", str(ROOT / "gaussianimage_b200" / "csrc"), f"OUT={out_dir}/libgi2d_trace.so", f"OBJ={out_dir}/obj",
                        "EXTRA=-DGI2D_TRACE"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
L = C.CDLL(str(out_dir / "libgi2d_trace.so"))
vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
L.gi2d_bin_fused_workspace_bytes.restype = sz
L.gi2d_bin_fused_workspace_bytes.argtypes = [i32] * 4
L.gi2d_bin_fused.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, i32, vp, sz, vp, vp, vp, vp, vp]
L.gi2d_raster_sum_fwd_packed.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
L.gi2d_raster_sum_bwd_packed.argtypes = [i32, i32, i32, i32, i32, vp, vp, vp, vp, vp]
L.gi2d_debug_set_trace.argtypes = [vp]
L.gi2d_raster_mse_workspace_bytes.restype = sz
L.gi2d_raster_sum_fwd_packed_mse.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, sz, vp]

args = [a for a in sys.argv[1:] if not a.startswith("--")]
cfg = args[0] if args else "C2"
s = make_config(cfg)
d = s.to("cuda:0")
dev = d.gt.device
H, W, (TX, TY, _) = s.H, s.W, s.tile_bounds
N = d.means2d.shape[0]
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
NW = 148 * 8 * 8
trace = torch.zeros(NW * 6, dtype=torch.int64, device=dev)
assert L.gi2d_debug_set_trace(trace.data_ptr()) == 0


def binning():
    assert L.gi2d_bin_fused(N, TX, TY, xys.data_ptr(), radii.data_ptr(), nth.data_ptr(), conics.data_ptr(), d.colors.data_ptr(),
                            d.opacity.data_ptr(), cap, ws.data_ptr(), ws_bytes, gids.data_ptr(), bins.data_ptr(), packed.data_ptr(),
                            status.data_ptr(), st) == 0


def fwd():
    assert L.gi2d_raster_sum_fwd_packed(TX, TY, W, H, bins.data_ptr(), packed.data_ptr(), status.data_ptr(), bg.data_ptr(),
                                        out.data_ptr(), 0, 0, st) == 0


mse_ws = torch.zeros(max(int(L.gi2d_raster_mse_workspace_bytes()), 4), dtype=torch.uint8, device=dev)
gt_chw = d.gt.permute(2, 0, 1).contiguous()
loss = torch.zeros(1, device=dev)


def fwd_loss():
    assert L.gi2d_raster_sum_fwd_packed_mse(TX, TY, W, H, bins.data_ptr(), packed.data_ptr(), status.data_ptr(), bg.data_ptr(),
                                            gt_chw.data_ptr(), 1, 1, v_out.data_ptr(), 0, loss.data_ptr(), mse_ws.data_ptr(),
                                            mse_ws.numel(), st) == 0


def bwd():
    assert L.gi2d_raster_sum_bwd_packed(N, TX, TY, W, H, bins.data_ptr(), packed.data_ptr(), v_out.data_ptr(), grads.data_ptr(), st) == 0


def pct(a, qs=(0, 10, 50, 90, 100)):
    return " / ".join(f"{np.percentile(a, q):.2f}" for q in qs)


GHZ = 1.965                                               # SM clock under load on this pool (bench.py's clocks line)
for warm in (True, False):
    for name, fn in (("raster_fwd", fwd), ("raster_fwd_loss", fwd_loss), ("raster_bwd", bwd)):
        for _ in range(3):
            binning(); fn()
        if warm:
            binning()
        else:
            flush.zero_()
        torch.cuda.synchronize()
        trace.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        t = trace.cpu().numpy().reshape(NW, 6)
        t = t[t[:, 1] != 0]
        work = t[:, 5] > 0
        sm = t[:, 4]
        entry_us = (t[:, 0] - t[:, 0].min()) / 1e3            # globaltimer, ns
        dur_us = (t[:, 3] - t[:, 1]) / GHZ / 1e3
        load_us = np.where(t[:, 2] > 0, (t[:, 2] - t[:, 1]) / GHZ / 1e3, 0.0)
        end_us = entry_us + dur_us
        print(f"== {cfg} {name} ({'input in L2 (just produced by the binning)' if warm else 'L2 flushed'}): event time "
              f"{1e3 * e0.elapsed_time(e1):.1f} us, {len(t)} warps, {int(work.sum())} with work, records/warp {pct(t[work, 5])}")
        print(f"   entry after the first warp's entry, us (min/p10/p50/p90/max): {pct(entry_us)}")
        print(f"   entry -> first chunk landed, us: {pct(load_us[work])}")
        print(f"   entry -> exit of warps with work, us: {pct(dur_us[work])}")
        print(f"   exit time of warps with work (after the first entry), us: {pct(end_us[work])};  last exit {end_us.max():.2f}")
        sms = np.unique(sm[work])
        per_sm_end = np.array([end_us[(sm == k) & work].max() for k in sms])
        per_sm_rec = np.array([t[(sm == k) & work, 5].sum() for k in sms])
        print(f"   per SM: last exit us {pct(per_sm_end)};  records {pct(per_sm_rec)};  SMs used {len(sms)}")
        print(f"   correlation(records of the SM, last exit of the SM) = {np.corrcoef(per_sm_rec, per_sm_end)[0, 1]:.2f}; "
              f"correlation(records of the warp, its duration) = {np.corrcoef(t[work, 5], dur_us[work])[0, 1]:.2f}")
        us_per_rec = (dur_us[work] - load_us[work]) / t[work, 5]
        print(f"   (duration - load latency) per record of the warp, ns: {pct(1e3 * us_per_rec)}")
        print("   " + "; ".join(f"at {100 * x:.0f} % ({x * end_us.max():.1f} us) {int(((entry_us <= x * end_us.max()) & (end_us > x * end_us.max()) & work).sum())} warps run"
                                for x in (0.25, 0.5, 0.75, 0.9)))
