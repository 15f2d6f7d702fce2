"""Adan with the constructor, state layout and step semantics of the reference's optimizer.Adan (optimizer.py:38-230), whose
step is ONE CUDA kernel per parameter group (csrc/adan.cu) instead of ~25 torch._foreach launches (optimizer.py:296-362).

    from gaussianimage_b200.optimizer import Adan
    opt = Adan(model.parameters(), lr=1e-3)                   # same arguments as the reference's class
    opt = Adan(model.parameters(), lr=1e-3, capturable=True)  # step() may be captured in a CUDA graph and replayed

State keys ('exp_avg', 'exp_avg_sq', 'exp_avg_diff', 'neg_pre_grad', group['step']) are the reference's, so
state_dict()s are interchangeable.  `foreach` / `fused` are accepted for signature compatibility; there is one
implementation and no CPU path (CPU parameters raise, as the reference's fused mode does, optimizer.py:224).

capturable=True keeps lr / betas / eps / weight decay and the step counter in device memory (gi2d_adan_multi_tensor_graph):
nothing in step() depends on host values that change from iteration to iteration, so forward + backward + step can be
one CUDA graph.  group['lr'] changes (LR schedulers) are noticed on the host and uploaded outside the graph by
`sync_hyperparameters()`; call it after scheduler.step() when replaying a graph.  max_grad_norm > 0 needs a host
read per step (optimizer.py:129-147) and cannot be combined with capturable.
"""
from __future__ import annotations

import ctypes as C
import math

import torch
from torch.optim.optimizer import Optimizer

from . import cuda as _C
from ._lib import check, lib
from .fused_adan import tables

_STATE_KEYS = ("exp_avg", "exp_avg_sq", "exp_avg_diff")


class Adan(Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0, max_grad_norm=0.0,
                 no_prox=False, foreach: bool = True, fused: bool = False, capturable: bool = False):
        # the reference's checks and messages (optimizer.py:78-93)
        if not 0.0 <= max_grad_norm:
            raise ValueError("Invalid Max grad norm: {}".format(max_grad_norm))
        if not 0.0 <= lr:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= eps:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        for i, b in enumerate(betas[:3]):
            if not 0.0 <= b < 1.0:
                raise ValueError("Invalid beta parameter at index {}: {}".format(i, b))
        if capturable and max_grad_norm > 0:
            raise ValueError("capturable=True cannot clip by the global gradient norm (it needs a host read per step)")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, max_grad_norm=max_grad_norm,
                                      no_prox=no_prox, foreach=foreach, fused=fused))
        self.capturable = bool(capturable)
        self._dev_state = {}          # id(group) -> (device block, host mirror tuple)

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault("no_prox", False)

    @torch.no_grad()
    def restart_opt(self):
        """optimizer.py:107-122: zero the moments and the step counter."""
        for group in self.param_groups:
            group["step"] = 0
            for p in group["params"]:
                if p.requires_grad:
                    st = self.state[p]
                    for k in _STATE_KEYS:
                        st[k] = torch.zeros_like(p)
        for gid in list(self._dev_state):
            self._dev_state.pop(gid)

    # ------------------------------------------------------------------------------------------------ helpers
    def _clip_factor(self) -> float:
        """optimizer.py:129-149."""
        if self.defaults["max_grad_norm"] <= 0:
            return 1.0
        dev = self.param_groups[0]["params"][0].device
        total = torch.zeros(1, device=dev)
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    total.add_(p.grad.pow(2).sum())
        norm = torch.sqrt(total)
        limit = torch.tensor(self.defaults["max_grad_norm"], device=dev)
        return torch.clamp(limit / (norm + self.param_groups[-1]["eps"]), max=1.0).item()

    def _hyper(self, group, clip):
        b1, b2, b3 = group["betas"]
        return (float(b1), float(b2), float(b3), float(group["lr"]), float(group["weight_decay"]), float(group["eps"]), float(clip))

    def _device_block(self, group, dev, clip, step_before):
        """Device-resident hyper-parameters of a group (capturable mode), re-uploaded when a host value changed."""
        L = lib()
        hyper = self._hyper(group, clip)
        entry = self._dev_state.get(id(group))
        if entry is None or entry[1] != hyper or entry[2] != dev:
            nbytes = L.gi2d_adan_state_bytes()
            host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            step = step_before if entry is None else None
            if step is None:           # keep the device counter: read it back (outside any capture)
                step = int(entry[0][nbytes - 8:nbytes - 4].view(torch.int32).item())
            check(L.gi2d_adan_state_init(host.data_ptr(), *hyper, int(step)), "adan_state_init")
            block = entry[0] if entry is not None and entry[2] == dev else torch.empty(nbytes, dtype=torch.uint8, device=dev)
            block.copy_(host, non_blocking=False)
            entry = (block, hyper, dev)
            self._dev_state[id(group)] = entry
        return entry[0]

    def sync_hyperparameters(self):
        """Upload changed lr / weight decay / betas of every group (call outside graph capture, e.g. after scheduler.step())."""
        for group in self.param_groups:
            entry = self._dev_state.get(id(group))
            if entry is not None and entry[1] != self._hyper(group, 1.0):
                self._device_block(group, entry[2], 1.0, None)

    def device_step(self, group_index: int = 0) -> int:
        """The step counter kept on the device (capturable mode)."""
        entry = self._dev_state.get(id(self.param_groups[group_index]))
        if entry is None:
            return int(self.param_groups[group_index].get("step", 0))
        n = entry[0].numel()
        return int(entry[0][n - 8:n - 4].view(torch.int32).item())

    # ------------------------------------------------------------------------------------------------ step
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        clip = self._clip_factor()
        L = lib()
        for group in self.param_groups:
            beta1, beta2, beta3 = group["betas"]
            capturing = self.capturable and torch.cuda.is_current_stream_capturing()
            step_before = group.get("step", 0)
            group["step"] = step = step_before + 1       # under replay the device counter is the truth (device_step())
            lists = [[], [], [], [], [], []]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    for k in _STATE_KEYS:
                        st[k] = torch.zeros_like(p)
                if "neg_pre_grad" not in st or step == 1:
                    st["neg_pre_grad"] = p.grad.clone().mul_(-clip)           # optimizer.py:186-188
                for lst, t in zip(lists, (p, p.grad, st["exp_avg"], st["exp_avg_sq"], st["exp_avg_diff"], st["neg_pre_grad"])):
                    lst.append(t)
            if not lists[0]:
                continue
            dev = lists[0][0].device
            if not dev.type == "cuda":
                raise ValueError("Fused Adan does not support CPU")
            n, nm, tb = tables(lists)
            with _C._guard(dev):
                stream = _C._stream(dev)
                if self.capturable:
                    if capturing and id(group) not in self._dev_state:
                        raise RuntimeError("Adan(capturable=True): run one step() outside the graph capture first "
                                           "(it creates the optimizer state and uploads the hyper-parameters)")
                    block = self._dev_state[id(group)][0] if capturing else self._device_block(group, dev, clip, step_before)
                    if n > 16:
                        raise RuntimeError("Adan(capturable=True) supports at most 16 tensors with gradients per group")
                    check(L.gi2d_adan_multi_tensor_graph(n, nm, *tb, block.data_ptr(), int(bool(group["no_prox"])), stream),
                          "adan_multi_tensor_graph")
                else:
                    check(L.gi2d_adan_multi_tensor(n, nm, *tb, float(beta1), float(beta2), float(beta3),
                                                   1.0 - beta1 ** step, 1.0 - beta2 ** step, math.sqrt(1.0 - beta3 ** step),
                                                   float(group["lr"]), float(group["weight_decay"]), float(group["eps"]),
                                                   int(bool(group["no_prox"])), float(clip), 1, stream), "adan_multi_tensor")
        return loss
