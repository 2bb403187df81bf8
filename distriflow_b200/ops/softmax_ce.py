"""Fused softmax-cross-entropy (forward + backward in one kernel launch on CUDA).

Standard form for the perf configs, dense-label form for the MNIST configs, and the reference's
swapped-argument quirk (`tf.losses.softmaxCrossEntropy(prediction, y)`,
`/root/reference/src/common/models.ts:139`) as `reference_compat`.
The CPU branch is the fp32 PyTorch restatement used as the numerics oracle.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import c_float, c_int, c_void_p, check, dt, ptr, stream

__all__ = ["softmax_cross_entropy", "reference_compat_ce"]


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_ce_sigs", False):
        L.dfb_softmax_ce_index.restype = c_int
        L.dfb_softmax_ce_index.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, ctypes.c_long,
                                           ctypes.c_long, c_int, c_float, ctypes.c_long, c_void_p]
        L.dfb_softmax_ce_dense.restype = c_int
        L.dfb_softmax_ce_dense.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, ctypes.c_long, ctypes.c_long,
                                           ctypes.c_long, c_int, c_float, c_int, c_void_p]
        L._dfb_ce_sigs = True
    return L


class _IndexCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index, inplace, unit_grad):
        L = _sigs()
        C = logits.shape[-1]
        x = logits.reshape(-1, C)
        if x.stride(1) != 1:
            x = x.contiguous()
        lab = labels.reshape(-1)
        if lab.dtype not in (torch.int64, torch.int32):
            lab = lab.long()
        lab = lab.contiguous()
        rows = x.shape[0]
        loss_rows = torch.empty(rows, dtype=torch.float32, device=x.device)
        need_grad = ctx.needs_input_grad[0]
        grad = None
        if need_grad:
            grad = x if inplace else torch.empty_like(x)
        n_valid = rows if ignore_index is None else int((lab != ignore_index).sum().item())
        n_valid = max(n_valid, 1)
        check(L.dfb_softmax_ce_index(ptr(x), ptr(lab), 0 if lab.dtype == torch.int64 else 1, ptr(loss_rows), ptr(grad), rows, C,
                                     ctypes.c_long(x.stride(0)), ctypes.c_long(grad.stride(0) if grad is not None else 0),
                                     dt(x), c_float(1.0 / n_valid), ctypes.c_long(-100 if ignore_index is None else ignore_index),
                                     stream()), "dfb_softmax_ce_index")
        ctx.unit_grad = unit_grad
        ctx.shape = logits.shape
        if need_grad:
            ctx.save_for_backward(grad)
        return loss_rows.sum() / n_valid

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        if not ctx.unit_grad:
            grad = grad * g.to(grad.dtype)
        return grad.reshape(ctx.shape), None, None, None, None


class _DenseCE(torch.autograd.Function):
    """mode 0: (z=logits*, y=labels); modes 1/2: reference-compat (z = one-hot target, y = prediction*); * = differentiated."""

    @staticmethod
    def forward(ctx, z, y, mode):
        L = _sigs()
        C = z.shape[-1]
        z2 = z.reshape(-1, C).contiguous()
        y2 = y.reshape(-1, C).to(z2.dtype).contiguous()
        rows = z2.shape[0]
        loss_rows = torch.empty(rows, dtype=torch.float32, device=z2.device)
        grad = torch.empty_like(z2)
        check(L.dfb_softmax_ce_dense(ptr(z2), ptr(y2), ptr(loss_rows), ptr(grad), rows, C, ctypes.c_long(C), ctypes.c_long(C),
                                     ctypes.c_long(C), dt(z2), c_float(1.0 / rows), mode, stream()), "dfb_softmax_ce_dense")
        ctx.save_for_backward(grad)
        ctx.mode = mode
        ctx.zshape, ctx.yshape = z.shape, y.shape
        return loss_rows.sum() / rows

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        grad = grad * g.to(grad.dtype)
        if ctx.mode == 0:
            return grad.reshape(ctx.zshape), None, None
        return None, grad.reshape(ctx.yshape), None


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: Optional[int] = None,
                          inplace: bool = False, unit_grad: bool = False) -> torch.Tensor:
    """Mean softmax-CE.  `labels`: class indices `[...]` or dense `[..., C]`.

    `inplace=True` lets the kernel overwrite `logits` with its gradient (saves one `[rows, C]` buffer; the
    caller must not use `logits` afterwards).  `unit_grad=True` promises the upstream gradient is 1 (the loss
    is the root of `backward()`), which skips one pass over the gradient."""
    index_labels = labels.dtype in (torch.int64, torch.int32) and labels.dim() == logits.dim() - 1
    if not logits.is_cuda:
        if index_labels:
            return F.cross_entropy(logits.float().reshape(-1, logits.shape[-1]), labels.reshape(-1).long(),
                                   ignore_index=-100 if ignore_index is None else ignore_index)
        return (-(labels.float() * F.log_softmax(logits.float(), dim=-1)).sum(-1)).mean()
    if index_labels:
        return _IndexCE.apply(logits, labels, ignore_index, inplace, unit_grad)
    return _DenseCE.apply(logits, labels, 0)


def reference_compat_ce(labels: torch.Tensor, preds: torch.Tensor, variant: str = "custom") -> torch.Tensor:
    """CUDA version of `losses.reference_compat_loss` (same `(labels, preds)` signature)."""
    return _DenseCE.apply(labels.to(preds.dtype), preds, 1 if variant == "custom" else 2)
