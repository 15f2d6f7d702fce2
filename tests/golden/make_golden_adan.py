"""Generates tests/golden/adan_reference.npz with the REFERENCE's own optimizer (imported from /root/reference, CPU):
trajectories of optimizer.Adan on three GaussianImage-shaped parameter tensors over 6 steps of seeded gradients, for the
default hyper-parameters, weight decay with and without no_prox, global-norm clipping, and the single-tensor path.
Run in the build container:  python tests/golden/make_golden_adan.py
"""
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from optimizer import Adan  # noqa: E402

SHAPES = [(97, 2), (97, 3), (97, 3), (5,)]
CASES = {
    "default": dict(lr=1e-3),
    "lr_model": dict(lr=0.018),                                 # the fitting scripts' learning rate (train.py --lr default)
    "wd": dict(lr=5e-3, weight_decay=0.02),
    "wd_no_prox": dict(lr=5e-3, weight_decay=0.02, no_prox=True),
    "clip": dict(lr=1e-3, max_grad_norm=0.5),
    "single": dict(lr=1e-3, weight_decay=0.01, foreach=False),
}
STEPS = 6


def main(out="tests/golden/adan_reference.npz"):
    data = {}
    for name, kw in CASES.items():
        g = torch.Generator().manual_seed(7)
        params = [torch.nn.Parameter(torch.randn(s, generator=g)) for s in SHAPES]
        opt = Adan(params, **kw)
        for i, p in enumerate(params):
            data[f"{name}_p0_{i}"] = p.detach().numpy().copy()
        for t in range(STEPS):
            for i, p in enumerate(params):
                scale = 1.0 if t % 2 == 0 else 3.0
                p.grad = torch.randn(p.shape, generator=g) * scale
                data[f"{name}_g{t}_{i}"] = p.grad.numpy().copy()
            opt.step()
            for i, p in enumerate(params):
                st = opt.state[p]
                data[f"{name}_p{t + 1}_{i}"] = p.detach().numpy().copy()
                if t + 1 not in (1, STEPS):
                    continue
                for k in ("exp_avg", "exp_avg_sq", "exp_avg_diff", "neg_pre_grad"):
                    data[f"{name}_{k}{t + 1}_{i}"] = st[k].numpy().copy()
    np.savez_compressed(out, **data)
    print("wrote", out, len(data), "arrays")


if __name__ == "__main__":
    main(*sys.argv[1:])
