"""CPU tests of the Adan oracle (oracle/adan_oracle.py) against trajectories of the REFERENCE's own optimizer.Adan
(tests/golden/adan_reference.npz, made by tests/golden/make_golden_adan.py from /root/reference/optimizer.py), and of the
host logic of gaussianimage_b200.optimizer.Adan that needs no GPU.

Tolerance: 2e-6 relative to the largest magnitude of the tensor per step (fp32; torch's CPU kernels fuse some
multiply-adds that numpy does not), written below.
"""
import numpy as np
import pytest

from oracle.adan_oracle import AdanOracle

GOLD = np.load("tests/golden/adan_reference.npz")
CASES = {
    "default": dict(lr=1e-3),
    "lr_model": dict(lr=0.018),
    "wd": dict(lr=5e-3, weight_decay=0.02),
    "wd_no_prox": dict(lr=5e-3, weight_decay=0.02, no_prox=True),
    "clip": dict(lr=1e-3, max_grad_norm=0.5),
    "single": dict(lr=1e-3, weight_decay=0.01),
}
NT, STEPS = 4, 6
RTOL = 2e-6


def close(got, want, what, rtol=RTOL):
    scale = np.abs(want).max() + 1e-30
    err = np.abs(got.astype(np.float64) - want.astype(np.float64)).max()
    assert err <= rtol * scale, f"{what}: max err {err:.3e} at scale {scale:.3e}"


@pytest.mark.parametrize("case", list(CASES))
def test_oracle_follows_the_reference_optimizer(case):
    o = AdanOracle([GOLD[f"{case}_p0_{i}"] for i in range(NT)], **CASES[case])
    for t in range(STEPS):
        o.step([GOLD[f"{case}_g{t}_{i}"] for i in range(NT)])
        for i in range(NT):
            close(o.p[i], GOLD[f"{case}_p{t + 1}_{i}"], f"{case} step {t + 1} param {i}", RTOL * (t + 1))
            if t + 1 in (1, STEPS):
                close(o.m[i], GOLD[f"{case}_exp_avg{t + 1}_{i}"], "exp_avg", RTOL * (t + 1))
                close(o.v[i], GOLD[f"{case}_exp_avg_sq{t + 1}_{i}"], "exp_avg_sq", RTOL * (t + 1))
                close(o.d[i], GOLD[f"{case}_exp_avg_diff{t + 1}_{i}"], "exp_avg_diff", RTOL * (t + 1))
                close(o.n[i], GOLD[f"{case}_neg_pre_grad{t + 1}_{i}"], "neg_pre_grad", RTOL * (t + 1))


def test_clip_case_really_clipped():
    o = AdanOracle([GOLD[f"clip_p0_{i}"] for i in range(NT)], **CASES["clip"])
    assert o.clip_factor([GOLD[f"clip_g0_{i}"] for i in range(NT)]) < 0.1


def test_constructor_checks_match_the_reference():
    """optimizer.py:78-93: same ValueErrors, same messages."""
    import torch
    from gaussianimage_b200.optimizer import Adan
    p = [torch.nn.Parameter(torch.zeros(3))]
    for kw, msg in ((dict(lr=-1.0), "Invalid learning rate: -1.0"), (dict(eps=-1.0), "Invalid epsilon value: -1.0"),
                    (dict(max_grad_norm=-1.0), "Invalid Max grad norm: -1.0"),
                    (dict(betas=(1.0, 0.9, 0.9)), "Invalid beta parameter at index 0: 1.0"),
                    (dict(betas=(0.9, 0.9, -0.1)), "Invalid beta parameter at index 2: -0.1")):
        with pytest.raises(ValueError, match=msg.replace(".", r"\.")):
            Adan(p, **kw)
    opt = Adan(p, lr=0.018)
    assert opt.defaults == dict(lr=0.018, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0, max_grad_norm=0.0,
                                no_prox=False, foreach=True, fused=False)
    p[0].grad = torch.ones(3)
    with pytest.raises(ValueError, match="Fused Adan does not support CPU"):     # optimizer.py:224: no CPU path
        opt.step()


def test_fused_adan_module_is_importable_under_the_name_the_reference_uses():
    import fused_adan            # optimizer.py:385
    assert callable(fused_adan.adan_multi_tensor) and callable(fused_adan.adan_single_tensor)
    import torch
    with pytest.raises(ValueError, match="Fused Adan does not support CPU"):
        t = [torch.zeros(4)]
        fused_adan.adan_multi_tensor(65536, None, [t, t, t, t, t, t], 0.98, 0.92, 0.99, 0.02, 0.08, 0.1, 1e-3, 0.0, 1e-8, False, 1.0)
