"""The `fused_adan` module the reference's optimizer imports when Adan(fused=True) (optimizer.py:385, :427, :453):
adan_multi_tensor / adan_single_tensor with the call signatures of optimizer.py:386-393 and :428-447, over
gi2d_adan_multi_tensor (include/gi2d.h, csrc/adan.cu).  The repository root has a top-level `fused_adan` package that
re-exports this module, so `PYTHONPATH=/root/repo` is all the reference needs.  No CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import torch
from torch import Tensor

from . import cuda as _C
from ._lib import check, lib

_VP16 = C.c_void_p * 16


def _table(tensors: Sequence[Tensor], name: str, numel: Sequence[int] | None):
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        if not t.is_cuda:
            raise ValueError("Fused Adan does not support CPU")            # optimizer.py:224, :231
        if t.dtype != torch.float32:
            raise RuntimeError(f"fused_adan: {name}[{i}] must be float32, got {t.dtype}")
        if not t.is_contiguous():
            raise RuntimeError(f"fused_adan: {name}[{i}] must be contiguous")
        if numel is not None and t.numel() != numel[i]:
            raise RuntimeError(f"fused_adan: {name}[{i}] has {t.numel()} elements, its parameter has {numel[i]}")
        arr[i] = t.data_ptr()
    return arr


def tables(tensor_lists: Sequence[Sequence[Tensor]]):
    """Host pointer tables of [params, grads, exp_avgs, exp_avg_sqs, exp_avg_diffs, neg_pre_grads]."""
    params = tensor_lists[0]
    n = len(params)
    if any(len(lst) != n for lst in tensor_lists) or len(tensor_lists) != 6:
        raise RuntimeError("fused_adan: expected six tensor lists of equal length")
    numel = [p.numel() for p in params]
    nm = (C.c_longlong * n)(*numel)
    names = ("params", "grads", "exp_avgs", "exp_avg_sqs", "exp_avg_diffs", "neg_pre_grads")
    return n, nm, [_table(lst, nme, numel) for lst, nme in zip(tensor_lists, names)]


def _launch(tensor_lists, beta1, beta2, beta3, bias_correction1, bias_correction2, bias_correction3_sqrt, lr, weight_decay,
            eps, no_prox, clip_global_grad_norm, update_neg):
    if len(tensor_lists[0]) == 0:
        return
    dev = tensor_lists[0][0].device
    if any(t.device != dev for lst in tensor_lists for t in lst):
        raise RuntimeError("fused_adan: all tensors of a call must be on one device")
    n, nm, tb = tables(tensor_lists)
    with _C._guard(dev):
        check(lib().gi2d_adan_multi_tensor(n, nm, *tb, float(beta1), float(beta2), float(beta3), float(bias_correction1),
                                           float(bias_correction2), float(bias_correction3_sqrt), float(lr),
                                           float(weight_decay), float(eps), int(bool(no_prox)),
                                           float(clip_global_grad_norm), int(update_neg), _C._stream(dev)), "adan_multi_tensor")


def adan_multi_tensor(chunk_size: int, noop_flag: Tensor, tensor_lists: List[List[Tensor]], beta1: float, beta2: float,
                      beta3: float, bias_correction1: float, bias_correction2: float, bias_correction3_sqrt: float, lr: float,
                      weight_decay: float, eps: float, no_prox: bool, clip_global_grad_norm) -> None:
    """optimizer.py:386-393.  chunk_size / noop_flag belong to apex-style multi_tensor_apply and are not needed: the
    kernel tiles the tensors itself.  neg_pre_grads is left to the caller (optimizer.py:394-395 rewrites it)."""
    _launch(tensor_lists, beta1, beta2, beta3, bias_correction1, bias_correction2, bias_correction3_sqrt, lr, weight_decay,
            eps, no_prox, clip_global_grad_norm, update_neg=0)


def adan_single_tensor(p_data_fp32: Tensor, out_p: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor,
                       exp_avg_diff: Tensor, neg_grad: Tensor, beta1: float, beta2: float, beta3: float,
                       bias_correction1: float, bias_correction2: float, bias_correction3_sqrt: float, lr: float,
                       weight_decay: float, eps: float, no_prox: bool, clip_global_grad_norm) -> None:
    """optimizer.py:428-447.  `p_data_fp32` is `param.data.float()` -- the parameter itself for fp32 parameters (the only
    kind the models have); for another dtype it is a temporary and the result is cast into `out_p`."""
    _launch([[p_data_fp32], [grad], [exp_avg], [exp_avg_sq], [exp_avg_diff], [neg_grad]], beta1, beta2, beta3,
            bias_correction1, bias_correction2, bias_correction3_sqrt, lr, weight_decay, eps, no_prox, clip_global_grad_norm,
            update_neg=0)
    if out_p.data_ptr() != p_data_fp32.data_ptr():
        out_p.copy_(p_data_fp32)
