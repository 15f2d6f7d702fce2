"""Adan step at the fitting size (C2: N = 70k Gaussians -> xyz (N,2), cholesky (N,3), colours (N,3) = 560k fp32 elements):
the reference's default foreach path (restated with torch._foreach ops, optimizer.py:296-362) against the one-kernel step,
eager and as CUDA-graph replays; plus the whole fit iteration (projection, rasterization, MSE, backward, Adan) as one graph.
python tools/bench_adan.py [--out gpurun_out/adan.json]"""
import argparse
import json
import math
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import gaussianimage_b200 as G
from gaussianimage_b200.optimizer import Adan
from gaussianimage_b200.synth import make_config

DEV = "cuda:0"


def foreach_step(P, Gs, M, V, D, Nn, step, lr=0.018, b1=0.98, b2=0.92, b3=0.99, eps=1e-8):
    bc1, bc2, bc3s = 1 - b1 ** step, 1 - b2 ** step, math.sqrt(1 - b3 ** step)
    torch._foreach_mul_(Gs, 1.0)
    torch._foreach_add_(Nn, Gs)
    torch._foreach_mul_(M, b1)
    torch._foreach_add_(M, Gs, alpha=1 - b1)
    torch._foreach_mul_(D, b2)
    torch._foreach_add_(D, Nn, alpha=1 - b2)
    torch._foreach_mul_(Nn, b2)
    torch._foreach_add_(Nn, Gs)
    torch._foreach_mul_(V, b3)
    torch._foreach_addcmul_(V, Nn, Nn, value=1 - b3)
    denom = torch._foreach_sqrt(V)
    torch._foreach_div_(denom, bc3s)
    torch._foreach_add_(denom, eps)
    torch._foreach_addcdiv_(P, M, denom, value=-(lr / bc1))
    torch._foreach_addcdiv_(P, D, denom, value=-(lr * b2 / bc2))
    torch._foreach_div_(P, 1.0)
    torch._foreach_zero_(Nn)
    torch._foreach_add_(Nn, Gs, alpha=-1.0)


def rate(fn, reps=300, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


def graphed(fn):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g.replay


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    N = 70_000
    shapes = [(N, 2), (N, 3), (N, 3)]
    mk = lambda: [torch.randn(s, device=DEV) for s in shapes]     # noqa: E731
    res = {"elements": sum(x * y for x, y in shapes), "bytes_per_step": 44 * sum(x * y for x, y in shapes)}
    P, Gs, M, V, D, Nn = mk(), mk(), mk(), [x.abs() for x in mk()], mk(), mk()
    cnt = [0]

    def fe():
        cnt[0] += 1
        foreach_step(P, Gs, M, V, D, Nn, cnt[0])
    res["foreach_eager_us"] = rate(fe)
    res["foreach_graph_us"] = rate(graphed(lambda: foreach_step(P, Gs, M, V, D, Nn, 100)))
    params = [torch.nn.Parameter(x) for x in mk()]
    for p, g in zip(params, Gs):
        p.grad = g
    opt = Adan(params, lr=0.018)
    res["fused_eager_us"] = rate(opt.step)
    optc = Adan(params, lr=0.018, capturable=True)
    optc.step()
    res["fused_graph_us"] = rate(graphed(optc.step))
    res["fused_graph_GBps"] = res["bytes_per_step"] / res["fused_graph_us"] / 1e3

    # whole fit iteration in one graph: projection + rasterization + MSE + backward + Adan
    s = make_config("C2")
    d = s.to(DEV)
    means = torch.nn.Parameter(torch.atanh(d.means2d.clamp(-0.999, 0.999)))
    chol = torch.nn.Parameter(d.cholesky.clone())
    cols = torch.nn.Parameter(d.colors.clone())
    G.set_sync_free(True)
    for capt, key in ((False, "foreach"), (True, "fused"), (True, "fused_and_loss_epilogue")):
        ps = [torch.nn.Parameter(x.detach().clone()) for x in (means, chol, cols)]
        o = Adan(ps, lr=0.018, capturable=True) if capt else None
        st = [[torch.zeros_like(p) for p in ps] for _ in range(4)]

        def it():
            for p in ps:
                p.grad = None
            xys, depths, radii, conics, nth = G.project_gaussians_2d(torch.tanh(ps[0]), ps[1], s.H, s.W, s.tile_bounds)
            if key == "fused_and_loss_epilogue":
                loss = G.rasterize_gaussians_sum_mse(xys, depths, radii, conics, nth, ps[2], d.opacity, s.H, s.W, d.gt, clamp=False)
            else:
                out = G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, ps[2], d.opacity, s.H, s.W)
                loss = torch.nn.functional.mse_loss(out, d.gt)
            loss.backward()
            if capt:
                o.step()
            else:
                with torch.no_grad():
                    foreach_step(ps, [p.grad for p in ps], st[0], st[1], st[2], st[3], 100)
        it()
        res[f"fit_iteration_with_{key}_adan_graph_us"] = rate(graphed(it), reps=300)
    G.check_overflow(DEV)
    res["speedup_step_graph"] = res["foreach_graph_us"] / res["fused_graph_us"]
    res["speedup_step_eager"] = res["foreach_eager_us"] / res["fused_eager_us"]
    txt = json.dumps(res, indent=1)
    print(txt)
    if a.out:
        Path(a.out).write_text(txt + "\n")


if __name__ == "__main__":
    main()
