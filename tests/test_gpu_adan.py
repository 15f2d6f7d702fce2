"""GPU tests of the fused Adan step (csrc/adan.cu through gaussianimage_b200.optimizer.Adan and the `fused_adan` seam)
against (a) the CPU oracle on the golden inputs of the reference's optimizer, (b) a torch fp32 restatement of the
reference's foreach path (optimizer.py:296-362) on the GPU at the fitting sizes.  Tolerance 2e-6 relative per step."""
import math

import numpy as np
import pytest
import torch

from gaussianimage_b200.optimizer import Adan
from oracle.adan_oracle import AdanOracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = np.load("tests/golden/adan_reference.npz")
CASES = {
    "default": dict(lr=1e-3),
    "lr_model": dict(lr=0.018),
    "wd": dict(lr=5e-3, weight_decay=0.02),
    "wd_no_prox": dict(lr=5e-3, weight_decay=0.02, no_prox=True),
    "clip": dict(lr=1e-3, max_grad_norm=0.5),
}
NT, STEPS, RTOL = 4, 6, 2e-6


def close(got, want, what, rtol):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    scale = np.abs(want).max() + 1e-30
    err = np.abs(got - want).max()
    assert err <= rtol * scale, f"{what}: max err {err:.3e} at scale {scale:.3e}"


@pytest.mark.parametrize("case", list(CASES))
def test_kernel_follows_oracle_and_reference_trajectory(case):
    params = [torch.nn.Parameter(torch.from_numpy(GOLD[f"{case}_p0_{i}"]).to(DEV)) for i in range(NT)]
    opt = Adan(params, **CASES[case])
    o = AdanOracle([GOLD[f"{case}_p0_{i}"] for i in range(NT)], **CASES[case])
    for t in range(STEPS):
        grads = [GOLD[f"{case}_g{t}_{i}"] for i in range(NT)]
        for p, g in zip(params, grads):
            p.grad = torch.from_numpy(g).to(DEV)
        opt.step()
        o.step(grads)
        for i, p in enumerate(params):
            st = opt.state[p]
            tol = RTOL * (t + 1)
            close(p.detach().cpu().numpy(), o.p[i], f"{case} step {t + 1} param {i} vs oracle", tol)
            close(p.detach().cpu().numpy(), GOLD[f"{case}_p{t + 1}_{i}"], f"{case} step {t + 1} param {i} vs reference", tol)
            close(st["exp_avg"].cpu().numpy(), o.m[i], "exp_avg", tol)
            close(st["exp_avg_sq"].cpu().numpy(), o.v[i], "exp_avg_sq", tol)
            close(st["exp_avg_diff"].cpu().numpy(), o.d[i], "exp_avg_diff", tol)
            close(st["neg_pre_grad"].cpu().numpy(), o.n[i], "neg_pre_grad", tol)
    assert opt.param_groups[0]["step"] == STEPS
    assert set(opt.state[params[0]]) == {"exp_avg", "exp_avg_sq", "exp_avg_diff", "neg_pre_grad"}


def _foreach_adan(params, grads, m, v, d, n, *, beta1, beta2, beta3, bc1, bc2, bc3s, lr, wd, eps, no_prox, clip):
    """torch restatement of the reference's default step (optimizer.py:296-362), test infrastructure."""
    torch._foreach_mul_(grads, clip)
    torch._foreach_add_(n, grads)
    torch._foreach_mul_(m, beta1)
    torch._foreach_add_(m, grads, alpha=1 - beta1)
    torch._foreach_mul_(d, beta2)
    torch._foreach_add_(d, n, alpha=1 - beta2)
    torch._foreach_mul_(n, beta2)
    torch._foreach_add_(n, grads)
    torch._foreach_mul_(v, beta3)
    torch._foreach_addcmul_(v, n, n, value=1 - beta3)
    denom = torch._foreach_sqrt(v)
    torch._foreach_div_(denom, bc3s)
    torch._foreach_add_(denom, eps)
    if no_prox:
        torch._foreach_mul_(params, 1 - lr * wd)
        torch._foreach_addcdiv_(params, m, denom, value=-(lr / bc1))
        torch._foreach_addcdiv_(params, d, denom, value=-(lr * beta2 / bc2))
    else:
        torch._foreach_addcdiv_(params, m, denom, value=-(lr / bc1))
        torch._foreach_addcdiv_(params, d, denom, value=-(lr * beta2 / bc2))
        torch._foreach_div_(params, 1 + lr * wd)
    torch._foreach_zero_(n)
    torch._foreach_add_(n, grads, alpha=-1.0)


@pytest.mark.parametrize("no_prox", [False, True])
def test_fused_adan_seam_matches_the_foreach_path_at_fitting_size(no_prox):
    """The call the reference makes with Adan(fused=True) (optimizer.py:383-395) against its default foreach path."""
    import fused_adan
    g = torch.Generator(device=DEV).manual_seed(3)
    shapes = [(70_000, 2), (70_000, 3), (70_000, 3), (7,), (1,)]
    mk = lambda: [torch.randn(s, generator=g, device=DEV) for s in shapes]      # noqa: E731
    P, M, V, D, Nn = mk(), mk(), [x.abs() for x in mk()], mk(), mk()
    ref = [[x.clone() for x in lst] for lst in (P, M, V, D, Nn)]
    b1, b2, b3, lr, wd, eps = 0.98, 0.92, 0.99, 0.018, 0.01, 1e-8
    for step in range(1, 5):
        G = mk()
        Gr = [x.clone() for x in G]
        kw = dict(beta1=b1, beta2=b2, beta3=b3, bc1=1 - b1 ** step, bc2=1 - b2 ** step, bc3s=math.sqrt(1 - b3 ** step), lr=lr,
                  wd=wd, eps=eps, no_prox=no_prox, clip=0.75)
        _foreach_adan(ref[0], Gr, ref[1], ref[2], ref[3], ref[4], **kw)
        # optimizer.py:383-395
        fused_adan.adan_multi_tensor(2048 * 32, torch.zeros(1, dtype=torch.int32, device=DEV), [P, G, M, V, D, Nn], b1, b2, b3,
                                     kw["bc1"], kw["bc2"], kw["bc3s"], lr, wd, eps, no_prox, 0.75)
        torch._foreach_zero_(Nn)
        torch._foreach_add_(Nn, G, alpha=-1.0)
        for name, ours, want in (("p", P, ref[0]), ("m", M, ref[1]), ("v", V, ref[2]), ("d", D, ref[3]), ("n", Nn, ref[4]), ("g", G, Gr)):
            for a, b in zip(ours, want):
                close(a.cpu().numpy(), b.cpu().numpy(), f"step {step} {name}", RTOL * step)


def test_single_tensor_entry_and_unaligned_views():
    import fused_adan
    g = torch.Generator(device=DEV).manual_seed(5)
    base = [torch.randn(1003, generator=g, device=DEV) for _ in range(6)]
    base[2].abs_()
    views = [b[1:1000] for b in base]                      # 4-byte aligned only: scalar path of the kernel
    want = [b.clone() for b in views]
    kw = dict(beta1=0.98, beta2=0.92, beta3=0.99, bc1=0.02, bc2=0.08, bc3s=0.1, lr=1e-2, wd=0.0, eps=1e-8, no_prox=False, clip=1.0)
    # order of `base`: p, g, v(abs), m, d, n  -> map explicitly
    p, gr, v, m, d, n = views
    wp, wg, wv, wm, wd_, wn = want
    _foreach_adan([wp], [wg], [wm], [wv], [wd_], [wn], **kw)
    fused_adan.adan_single_tensor(p, p, gr, m, v, d, n, 0.98, 0.92, 0.99, 0.02, 0.08, 0.1, 1e-2, 0.0, 1e-8, False, 1.0)
    n.zero_().add_(gr, alpha=-1.0)                         # optimizer.py:449
    for a, b, nm in ((p, wp, "p"), (m, wm, "m"), (v, wv, "v"), (d, wd_, "d"), (n, wn, "n")):
        close(a.cpu().numpy(), b.cpu().numpy(), nm, RTOL)
    for b in base:                                         # nothing outside the views was touched
        assert torch.isfinite(b).all()
    assert torch.equal(base[0][:1], base[0][:1]) and base[0].numel() == 1003


def test_capturable_step_replays_as_successive_steps():
    g = torch.Generator(device=DEV).manual_seed(11)
    shapes = [(5000, 2), (5000, 3), (5000, 3)]
    init = [torch.randn(s, generator=g, device=DEV) for s in shapes]
    grads = [[torch.randn(s, generator=g, device=DEV) for s in shapes] for _ in range(6)]
    # eager trajectory
    pe = [torch.nn.Parameter(x.clone()) for x in init]
    oe = Adan(pe, lr=0.018, weight_decay=0.01)
    for t in range(6):
        for p, gr in zip(pe, grads[t]):
            p.grad = gr.clone()
        if t == 4:
            oe.param_groups[0]["lr"] = 0.009       # an LR-scheduler step
        oe.step()
    # graph trajectory: step 1 eager (creates the state), steps 2..6 are replays of ONE captured step
    pg = [torch.nn.Parameter(x.clone()) for x in init]
    og = Adan(pg, lr=0.018, weight_decay=0.01, capturable=True)
    static_g = [torch.zeros(s, device=DEV) for s in shapes]
    for p, sg in zip(pg, static_g):
        p.grad = sg
    for sg, gr in zip(static_g, grads[0]):
        sg.copy_(gr)
    og.step()
    assert og.device_step() == 1
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        og.step()
    # capture does not execute: roll the host counter back to what has really run
    for t in range(1, 6):
        for sg, gr in zip(static_g, grads[t]):
            sg.copy_(gr)
        if t == 4:
            og.param_groups[0]["lr"] = 0.009
            og.sync_hyperparameters()
        graph.replay()
    torch.cuda.synchronize()
    assert og.device_step() == 6
    # device-side pow() vs Python's: the bias corrections may differ in the last bit of a double, so compare to 1e-6
    for a, b in zip(pg, pe):
        close(a.detach().cpu().numpy(), b.detach().cpu().numpy(), "graph-replayed steps vs eager steps", 1e-6)
        for k in ("exp_avg", "exp_avg_sq", "exp_avg_diff", "neg_pre_grad"):
            close(og.state[a][k].cpu().numpy(), oe.state[b][k].cpu().numpy(), k, 1e-6)
