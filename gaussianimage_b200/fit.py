"""The reference's per-image training step as ONE CUDA graph -- measurement / test helper, not a re-implementation of the
reference's trainer (train.py's argument parsing, logging, checkpoints and the model classes stay the caller's).

What one replay does is what `GaussianImage_Cholesky.train_iter` does per iteration (gaussianimage_cholesky.py:100-125 with
forward :76-85): means = tanh(_xyz), L = _cholesky + cholesky_bound, project_gaussians_2d, rasterize_gaussians_sum,
clamp to [0,1], L2 loss against the (1,3,H,W) target, backward, Adan step (optimizer.py, lr from the StepLR schedule of
gaussianimage_cholesky.py:55: step 20000, gamma 0.5).  Parameters are initialised like the model's __init__ (:30-34).

    fused=True   loss epilogue inside the rasterizer (loss.py) + one-kernel capturable Adan (optimizer.py)
    fused=False  the reference's own sequence through the operator API: rasterize_gaussians_sum, torch.clamp, permute,
                 F.mse_loss, and the foreach Adan arithmetic of optimizer.py:296-362 restated with torch._foreach ops
                 (bias corrections frozen per capture, so only usable for short comparisons)
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from . import project_gaussians_2d, rasterize_gaussians_sum, rasterize_gaussians_sum_mse
from .optimizer import Adan


class GraphFitter:
    def __init__(self, gt_chw: torch.Tensor, num_points: int, lr: float = 0.018, seed: int = 1, fused: bool = True,
                 lr_step: int = 20000, lr_gamma: float = 0.5, check_every: int = 256):
        assert gt_chw.dim() == 4 and gt_chw.shape[:2] == (1, 3) and gt_chw.is_cuda
        dev = gt_chw.device
        self.H, self.W = int(gt_chw.shape[2]), int(gt_chw.shape[3])
        self.tile_bounds = ((self.W + 15) // 16, (self.H + 15) // 16, 1)
        self.gt, self.fused = gt_chw.contiguous(), fused
        g = torch.Generator().manual_seed(seed)
        N = num_points
        # gaussianimage_cholesky.py:30-34
        self._xyz = torch.nn.Parameter(torch.atanh(2 * (torch.rand(N, 2, generator=g) - 0.5)).to(dev))
        self._cholesky = torch.nn.Parameter(torch.rand(N, 3, generator=g).to(dev))
        self._features_dc = torch.nn.Parameter(torch.rand(N, 3, generator=g).to(dev))
        self._opacity = torch.ones(N, 1, device=dev)
        self.cholesky_bound = torch.tensor([0.5, 0.0, 0.5], device=dev).view(1, 3)
        self.params = [self._xyz, self._cholesky, self._features_dc]
        self.lr0, self.lr_step, self.lr_gamma = lr, lr_step, lr_gamma
        self.iteration = 0
        self.opt = Adan(self.params, lr=lr, capturable=True) if fused else None
        self._state = None if fused else [[torch.zeros_like(p) for p in self.params] for _ in range(4)]   # m, v, d, neg_pre_grad
        self.loss = None
        self.graph = None
        self.check_every = check_every      # replays between two reads of the sticky overflow status (0 = never)
        self.recaptures = 0

    # one training iteration, eager
    def step(self):
        for p in self.params:
            p.grad = None
        means = torch.tanh(self._xyz)                                            # :62
        L = self._cholesky + self.cholesky_bound                                 # :74
        pr = project_gaussians_2d(means, L, self.H, self.W, self.tile_bounds)    # :77
        if self.fused:
            loss = rasterize_gaussians_sum_mse(*pr, self._features_dc, self._opacity, self.H, self.W, self.gt, clamp=True)
        else:
            out = rasterize_gaussians_sum(*pr, self._features_dc, self._opacity, self.H, self.W)           # :80-82
            image = torch.clamp(out, 0, 1).view(-1, self.H, self.W, 3).permute(0, 3, 1, 2).contiguous()      # :84-85
            loss = F.mse_loss(image, self.gt)                                                              # utils.py:21
        loss.backward()
        if self.fused:
            self.opt.step()
        else:
            self._foreach_adan()
        self.loss = loss.detach()
        return self.loss

    def _foreach_adan(self, b1=0.98, b2=0.92, b3=0.99, eps=1e-8):
        """optimizer.py:296-362 with the step count of the moment of the call (frozen inside a captured graph)."""
        t = self.iteration + 1
        lr = self.lr0 * self.lr_gamma ** (self.iteration // self.lr_step)
        grads = [p.grad for p in self.params]
        m, v, d, n = self._state
        with torch.no_grad():
            if t == 1:
                torch._foreach_zero_(n)
                torch._foreach_add_(n, grads, alpha=-1.0)
            torch._foreach_add_(n, grads)
            torch._foreach_mul_(m, b1)
            torch._foreach_add_(m, grads, alpha=1 - b1)
            torch._foreach_mul_(d, b2)
            torch._foreach_add_(d, n, alpha=1 - b2)
            torch._foreach_mul_(n, b2)
            torch._foreach_add_(n, grads)
            torch._foreach_mul_(v, b3)
            torch._foreach_addcmul_(v, n, n, value=1 - b3)
            denom = torch._foreach_sqrt(v)
            torch._foreach_div_(denom, math.sqrt(1 - b3 ** t))
            torch._foreach_add_(denom, eps)
            torch._foreach_addcdiv_(self.params, m, denom, value=-(lr / (1 - b1 ** t)))
            torch._foreach_addcdiv_(self.params, d, denom, value=-(lr * b2 / (1 - b2 ** t)))
            torch._foreach_zero_(n)
            torch._foreach_add_(n, grads, alpha=-1.0)

    def capture(self):
        """First iteration eagerly (creates optimizer state, learns the intersection capacity), then one captured."""
        assert self.fused, "only the fused step has nothing that changes on the host from iteration to iteration"
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self.step()
            self.iteration += 1
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        from . import graphs as GR
        g = GR.new_graph()
        with torch.cuda.graph(g):
            self.step()
        GR.fold_memsets(g)              # memset nodes (none on the fused path; ATen's reduction semaphores otherwise) -> fill kernels
        self.graph = g
        return g

    def run(self, iterations: int):
        """`iterations` more training iterations (graph replays when captured); applies the StepLR schedule."""
        for _ in range(iterations):
            if self.fused and self.iteration and self.iteration % self.lr_step == 0:
                self.opt.param_groups[0]["lr"] = self.lr0 * self.lr_gamma ** (self.iteration // self.lr_step)
                self.opt.sync_hyperparameters()
            if self.graph is not None:
                if self.check_every and self.iteration % self.check_every == 0:
                    self._check_overflow()
                self.graph.replay()
            else:
                self.step()
            self.iteration += 1
        if self.graph is not None and self.check_every:
            self._check_overflow()
        return self.loss

    def _check_overflow(self):
        """A replayed graph cannot grow its workspace: as the Gaussians spread, the intersection count may outgrow the capacity
        frozen at capture time.  The binning chain reports that in a sticky device word; read it now and then, and on
        overflow -- the capacity has been raised by check_overflow() -- capture the step again.  (The replays since the
        overflow saw truncated tile lists: at most `check_every` slightly wrong gradient steps, then exact again.)"""
        from .rasterize_sum import check_overflow
        try:
            check_overflow(self.gt.device)
        except RuntimeError:
            self.recaptures += 1
            self.graph = None
            self.capture()

    def psnr(self) -> float:
        """gaussianimage_cholesky.py:118-119 (for the L2 loss, mse_loss(image, gt) is the loss itself)."""
        return 10 * math.log10(1.0 / float(self.loss.item()))
