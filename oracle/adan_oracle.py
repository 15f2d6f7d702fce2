"""TEST INFRASTRUCTURE -- CPU restatement (numpy, fp32 element arithmetic, Python-double scalars) of the reference's Adan
step: optimizer.py:124-230 (`Adan.step`) and :232-279 (`_single_tensor_adan`, the arithmetic definition that the foreach
and fused variants must reproduce).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.

Parity pinning status: PINNED -- tests/golden/adan_reference.npz holds trajectories produced by the reference's own
optimizer.Adan imported from /root/reference (tests/golden/make_golden_adan.py, CPU, foreach and single-tensor paths);
tests/test_adan_cpu.py checks this file against them.
"""
import math

import numpy as np

f32 = np.float32


class AdanOracle:
    """State and step of optimizer.Adan for a list of fp32 arrays (one parameter group)."""

    def __init__(self, params, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0, max_grad_norm=0.0, no_prox=False):
        self.p = [np.array(p, dtype=f32) for p in params]
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.max_grad_norm, self.no_prox = max_grad_norm, no_prox
        self.step_count = 0
        self.m = [np.zeros_like(p) for p in self.p]         # exp_avg       (optimizer.py:181)
        self.v = [np.zeros_like(p) for p in self.p]         # exp_avg_sq    (:182)
        self.d = [np.zeros_like(p) for p in self.p]         # exp_avg_diff  (:183)
        self.n = [None] * len(self.p)                       # neg_pre_grad  (:186-188)

    def clip_factor(self, grads):
        """optimizer.py:129-149 (fp32 tensor arithmetic, then .item())."""
        if self.max_grad_norm <= 0:
            return 1.0
        total = f32(0)
        for g in grads:
            total = f32(total + np.sum(np.square(g, dtype=f32), dtype=f32))
        norm = np.sqrt(total, dtype=f32)
        return float(min(f32(f32(self.max_grad_norm) / f32(norm + f32(self.eps))), f32(1.0)))

    def step(self, grads):
        """grads are scaled in place by the clip factor, as optimizer.py:240 does."""
        grads = [np.asarray(g, dtype=f32) for g in grads]
        clip = self.clip_factor(grads)
        b1, b2, b3 = self.betas
        self.step_count += 1
        t = self.step_count
        bc1, bc2, bc3s = 1.0 - b1 ** t, 1.0 - b2 ** t, math.sqrt(1.0 - b3 ** t)       # :169-171, :212
        step_size, step_size_diff = self.lr / bc1, self.lr * b2 / bc2                 # :256-257
        out_g = []
        for i, g in enumerate(grads):
            if self.n[i] is None or t == 1:
                self.n[i] = (g * f32(-clip)).astype(f32)                               # :186-188
            g = (g * f32(clip)).astype(f32)                                            # :240
            n = (self.n[i] + g).astype(f32)                                            # :244
            self.m[i] = (self.m[i] * f32(b1)).astype(f32) + f32(1 - b1) * g            # :246
            self.d[i] = (self.d[i] * f32(b2)).astype(f32) + f32(1 - b2) * n            # :247
            n = (n * f32(b2)).astype(f32) + g                                          # :250
            self.v[i] = (self.v[i] * f32(b3)).astype(f32) + f32(1 - b3) * n * n        # :251
            denom = np.sqrt(self.v[i], dtype=f32) / f32(bc3s) + f32(self.eps)          # :255
            p = self.p[i]
            if self.no_prox:
                p = p * f32(1 - self.lr * self.wd)                                     # :260
                p = p + f32(-step_size) * (self.m[i] / denom)                          # :261
                p = p + f32(-step_size_diff) * (self.d[i] / denom)
            else:
                p = p + f32(-step_size) * (self.m[i] / denom)                          # :264
                p = p + f32(-step_size_diff) * (self.d[i] / denom)
                p = p / f32(1 + self.lr * self.wd)                                     # :266
            self.p[i] = p.astype(f32)
            self.n[i] = (-g).astype(f32)                                               # :268
            out_g.append(g)
        return out_g
