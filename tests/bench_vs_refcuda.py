"""TEST/MEASUREMENT INFRASTRUCTURE (lives under tests/ because it executes oracle/_ref): times the REFERENCE's own CUDA
kernels (oracle/_ref, compiled from /root/reference by oracle/Makefile) around the reference's own Python staging
(gsplat/gsplat/utils.py:116-167, rasterize_sum.py:109-185: torch.zeros outputs, cumsum + .item(), torch.sort, gather)
beside this repository's operator API on the same synthetic scene, same box, same process.

BASELINE.json's target is ">= the reference gsplat CUDA build's render FPS and fit iters/s at 768x512 with 70k
Gaussians on one B200"; no such number is published, so it is measured here.  Not part of bench.py: the
product and its bench never load oracle/.

    python tests/bench_vs_refcuda.py [--config C2] [--reps 200] [--out gpurun_out/vs_refcuda_C2.json]

Timing: wall clock around `reps` back-to-back steps with a device synchronize on both sides (the reference's step contains
a host read, so CUDA-event time and wall time coincide for it); L2 is not flushed (both sides equally warm).
"""
import argparse
import ctypes as C
import json
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "tests")]
import refcuda  # noqa: E402
import gaussianimage_b200 as G  # noqa: E402
from gaussianimage_b200.synth import CONFIGS, make_scene  # noqa: E402


def _p(t):
    return C.c_void_p(t.data_ptr())


class RefPipeline:
    """The reference operator sequence, allocation for allocation (bindings.cu torch::zeros, utils.py staging)."""

    def __init__(self, d, s, flavor, param):
        self.d, self.s, self.L, self.param = d, s, refcuda.lib(flavor), param
        self.bg = torch.ones(3)          # host: passed to the kernels by value
        self.alt_colors = d.colors.flip(0).contiguous() * 0.5
        self.one_colors = torch.ones_like(d.colors)

    def project(self):
        d, s, dev = self.d, self.s, self.d.means2d.device
        N = d.means2d.shape[0]
        TX, TY = s.tile_bounds[:2]
        xys, depths = torch.zeros(N, 2, device=dev), torch.zeros(N, device=dev)
        radii = torch.zeros(N, dtype=torch.int32, device=dev)
        conics, nth = torch.zeros(N, 3, device=dev), torch.zeros(N, dtype=torch.int32, device=dev)
        if self.param == "rs":
            self.L.ref_project_scale_rot_fwd(N, _p(d.means2d), _p(d.scales), _p(d.rotation), s.H, s.W, TX, TY, C.c_float(0.01), _p(xys),
                                             _p(depths), _p(radii), _p(conics), _p(nth))
        else:
            self.L.ref_project_cholesky_fwd(N, _p(d.means2d), _p(d.cholesky), s.H, s.W, TX, TY, C.c_float(0.01), _p(xys), _p(depths),
                                            _p(radii), _p(conics), _p(nth))
        return xys, depths, radii, conics, nth

    def model_forward(self):
        """gaussianimage_cholesky.py:76-98: one projection, three rasterize_gaussians_sum calls (Q13)."""
        pr = self.project()
        c = self.d.colors
        return [self.rasterize(*pr, colors=k)[0] for k in (c, self.alt_colors, self.one_colors)]

    def rasterize(self, xys, depths, radii, conics, nth, colors=None):
        d, s, dev = self.d, self.s, self.d.means2d.device
        colors = d.colors if colors is None else colors
        N, H, W = xys.shape[0], s.H, s.W
        TX, TY = s.tile_bounds[:2]
        cum = torch.cumsum(nth, dim=0, dtype=torch.int32)                    # utils.py:116
        n = int(cum[-1].item())                                              # utils.py:117 (host sync, Q12)
        ids = torch.zeros(n, dtype=torch.int64, device=dev)                  # bindings.cu:289-296
        gids = torch.zeros(n, dtype=torch.int32, device=dev)
        self.L.ref_map_gaussian_to_intersects(N, _p(xys), _p(depths), _p(radii), _p(cum), TX, TY, _p(ids), _p(gids))
        ids_s, order = torch.sort(ids)                                       # utils.py:164
        gids_s = torch.gather(gids, 0, order)                                # utils.py:165
        bins = torch.zeros(max(n, TX * TY), 2, dtype=torch.int32, device=dev)   # bindings.cu:319-321 (Q4 made safe)
        self.L.ref_get_tile_bin_edges(n, _p(ids_s), _p(bins))
        out, Ts = torch.zeros(H, W, 3, device=dev), torch.zeros(H, W, device=dev)   # bindings.cu:443-451
        fidx = torch.zeros(H, W, dtype=torch.int32, device=dev)
        self.L.ref_rasterize_sum_fwd(TX, TY, W, H, _p(gids_s), _p(bins), _p(xys), _p(conics), _p(colors), _p(d.opacity), _p(self.bg),
                                     _p(out), _p(Ts), _p(fidx))
        return out, (gids_s, bins, Ts, fidx)

    def render(self):
        return self.rasterize(*self.project())[0]

    def fit(self):
        d, s, dev = self.d, self.s, self.d.means2d.device
        H, W = s.H, s.W
        TX, TY = s.tile_bounds[:2]
        xys, depths, radii, conics, nth = self.project()
        N = xys.shape[0]
        out, (gids_s, bins, Ts, fidx) = self.rasterize(xys, depths, radii, conics, nth)
        v_out = (2.0 / out.numel()) * (out - d.gt)                           # gradient of F.mse_loss
        v_alpha = torch.zeros(H, W, device=dev)
        v_xy, v_conic = torch.zeros(N, 2, device=dev), torch.zeros(N, 3, device=dev)      # bindings.cu:752-756
        v_col, v_op = torch.zeros(N, 3, device=dev), torch.zeros(N, 1, device=dev)
        self.L.ref_rasterize_sum_bwd(TX, TY, W, H, _p(gids_s), _p(bins), _p(xys), _p(conics), _p(d.colors), _p(d.opacity), _p(self.bg),
                                     _p(Ts), _p(fidx), _p(v_out), _p(v_alpha), _p(v_xy), _p(v_conic), _p(v_col), _p(v_op))
        v_cov, v_mean = torch.zeros(N, 3, device=dev), torch.zeros(N, 2, device=dev)       # bindings.cu:919-928
        v_depth = torch.zeros(N, device=dev)
        if self.param == "rs":
            v_scale, v_rot = torch.zeros(N, 2, device=dev), torch.zeros(N, 1, device=dev)
            self.L.ref_project_scale_rot_bwd(N, _p(d.means2d), _p(d.scales), _p(d.rotation), H, W, _p(radii), _p(conics), _p(v_xy),
                                             _p(v_depth), _p(v_conic), _p(v_cov), _p(v_mean), _p(v_scale), _p(v_rot))
            return v_mean, v_scale, v_rot, v_col
        v_L = torch.zeros(N, 3, device=dev)
        self.L.ref_project_cholesky_bwd(N, _p(d.means2d), _p(d.cholesky), H, W, _p(radii), _p(conics), _p(v_xy), _p(v_depth), _p(v_conic),
                                        _p(v_cov), _p(v_mean), _p(v_L))
        return v_mean, v_L, v_col


class OurPipeline:
    def __init__(self, d, s, param):
        self.d, self.s, self.param = d, s, param
        self.means = d.means2d.clone().requires_grad_(True)
        self.colors = d.colors.clone().requires_grad_(True)
        self.alt_colors = d.colors.flip(0).contiguous() * 0.5
        self.one_colors = torch.ones_like(d.colors)
        self.p = ([d.scales.clone().requires_grad_(True), d.rotation.clone().requires_grad_(True)] if param == "rs"
                  else [d.cholesky.clone().requires_grad_(True)])

    def forward(self):
        s = self.s
        if self.param == "rs":
            pr = G.project_gaussians_2d_scale_rot(self.means, self.p[0], self.p[1], s.H, s.W, s.tile_bounds)
        else:
            pr = G.project_gaussians_2d(self.means, self.p[0], s.H, s.W, s.tile_bounds)
        return G.rasterize_gaussians_sum(*pr, self.colors, self.d.opacity, s.H, s.W)

    def render(self):
        with torch.no_grad():
            return self.forward()

    def model_forward(self):
        s, d = self.s, self.d
        with torch.no_grad():
            if self.param == "rs":
                pr = G.project_gaussians_2d_scale_rot(self.means, self.p[0], self.p[1], s.H, s.W, s.tile_bounds)
            else:
                pr = G.project_gaussians_2d(self.means, self.p[0], s.H, s.W, s.tile_bounds)
            return [G.rasterize_gaussians_sum(*pr, k, d.opacity, s.H, s.W) for k in (self.colors, self.alt_colors, self.one_colors)]

    def fit(self):
        loss = torch.nn.functional.mse_loss(self.forward(), self.d.gt)
        return torch.autograd.grad(loss, [self.means, *self.p, self.colors])


def rate(fn, reps, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return reps / (time.perf_counter() - t0)


def graph_rate(fn, reps):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        keep = fn()      # noqa: F841  (outputs stay alive with the graph)
    return rate(g.replay, reps)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    H, W, N, param = CONFIGS[a.config]
    s = make_scene(H, W, N, seed=1)
    d = s.to("cuda:0")
    res = {"config": a.config, "H": H, "W": W, "N": N, "param": param, "reps": a.reps, "device": torch.cuda.get_device_name(0),
           "timing": "wall clock over back-to-back steps, synchronize on both sides, warm L2, legacy default stream"}
    reps = a.reps if N <= 200_000 else max(20, a.reps // 10)
    for flavor in ("fast", "jit"):
        if not refcuda.available(flavor):
            continue
        R = RefPipeline(d, s, flavor, param)
        res[f"reference_cuda_{flavor}"] = {"render_fps": rate(R.render, reps), "fit_iters_per_s": rate(R.fit, reps),
                                           "model_forward_per_s": rate(R.model_forward, reps),
                                           "flags": "-O3 --use_fast_math (pip build, setup.py:74-86)" if flavor == "fast"
                                           else "-O3 (JIT build, _backend.py:39-40)"}
    O = OurPipeline(d, s, param)
    ours = {"render_fps_eager_strict": rate(O.render, reps), "fit_iters_per_s_eager_strict": rate(O.fit, reps)}
    ours["model_forward_per_s_eager_strict"] = rate(O.model_forward, reps)
    G.set_memoise_binning(False)
    ours["model_forward_per_s_eager_strict_no_memo"] = rate(O.model_forward, reps)
    G.set_memoise_binning(True)
    G.set_sync_free(True)
    ours["model_forward_per_s_graph"] = graph_rate(O.model_forward, reps)
    ours["render_fps_eager_sync_free"] = rate(O.render, reps)
    ours["fit_iters_per_s_eager_sync_free"] = rate(O.fit, reps)
    ours["render_fps_graph"] = graph_rate(O.render, reps)
    ours["fit_iters_per_s_graph"] = graph_rate(O.fit, reps)
    G.check_overflow("cuda:0")
    G.set_sync_free(False)
    res["gi2d_b200"] = ours
    ref = res.get("reference_cuda_fast") or res.get("reference_cuda_jit")
    if ref:
        res["speedup_vs_reference_cuda"] = {
            "render_eager_strict": ours["render_fps_eager_strict"] / ref["render_fps"],
            "render_graph": ours["render_fps_graph"] / ref["render_fps"],
            "fit_eager_strict": ours["fit_iters_per_s_eager_strict"] / ref["fit_iters_per_s"],
            "fit_graph": ours["fit_iters_per_s_graph"] / ref["fit_iters_per_s"],
            "model_forward_eager_strict": ours["model_forward_per_s_eager_strict"] / ref["model_forward_per_s"],
            "model_forward_graph": ours["model_forward_per_s_graph"] / ref["model_forward_per_s"],
        }
    txt = json.dumps(res, indent=1)
    print(txt)
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(txt + "\n")


if __name__ == "__main__":
    main()
