"""Fused residual-add + LayerNorm on CUDA (`csrc/norm.cu`).

`layer_norm(x, gamma, beta)` -> y, and `add_layer_norm(x, r, gamma, beta)` -> (s = x + r, y = LN(s)): the transformer
residual stream and its normalisation in one pass forward and one pass backward (the backward also folds the gradient
arriving through the residual stream into dx).  Parameter gradients are reduced deterministically (per-block partials
+ fixed-order reduce) and, under the engine, written straight into the flat gradient bucket.
"""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import c_float, c_int, c_void_p, check, dt, ptr, stream


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_norm_sigs", False):
        L.dfb_add_layernorm_fwd.restype = c_int
        L.dfb_add_layernorm_fwd.argtypes = [c_void_p] * 8 + [c_int, c_int, c_float, c_int, c_void_p]
        L.dfb_layernorm_bwd_blocks.restype = c_int
        L.dfb_layernorm_bwd_blocks.argtypes = [c_int]
        L.dfb_layernorm_bwd_staged.restype = c_int
        L.dfb_layernorm_bwd_staged.argtypes = [c_int]
        L.dfb_add_layernorm_bwd.restype = c_int
        L.dfb_add_layernorm_bwd.argtypes = [c_void_p] * 11 + [c_int, c_int, c_int, c_void_p]
        L.dfb_permute_rows.restype = c_int
        L.dfb_permute_rows.argtypes = [c_void_p, c_void_p] + [c_int] * 4 + [ctypes.c_long] * 8 + [c_int, c_void_p]
        L.dfb_bn_partial_blocks.restype = c_int
        L.dfb_bn_partial_blocks.argtypes = [ctypes.c_long, c_int, c_int]
        L.dfb_bn_fwd_train.restype = c_int
        L.dfb_bn_fwd_train.argtypes = [c_void_p] * 11 + [ctypes.c_long, c_int, c_float, c_float, c_int, c_int, c_int, c_void_p]
        L.dfb_bn_fwd_eval.restype = c_int
        L.dfb_bn_fwd_eval.argtypes = [c_void_p] * 8 + [ctypes.c_long, c_int, c_float, c_int, c_int, c_int, c_void_p]
        L.dfb_bn_bwd.restype = c_int
        L.dfb_bn_bwd.argtypes = [c_void_p] * 12 + [ctypes.c_long, c_int, c_int, c_int, c_int, c_void_p]
        L._dfb_norm_sigs = True
    return L


import os as _os

# bias gradients produced by the kernel that produces dY (LayerNorm backward, attention backward) instead of a separate
# column-sum pass over dY; DFB_FUSE_BIAS_GRAD=0 restores the separate kernel for A/B runs
FUSE_BIAS_GRAD = _os.environ.get("DFB_FUSE_BIAS_GRAD", "1") == "1"


def _supported(x: torch.Tensor) -> bool:
    D = x.shape[-1]
    vn = 8 if x.dtype == torch.bfloat16 else 4
    return x.dtype in (torch.bfloat16, torch.float32) and D % vn == 0 and D <= 32 * vn * 8


class _AddLN(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, r, gamma, beta, eps):
        L = _sigs()
        D = x.shape[-1]
        x2 = x.reshape(-1, D).contiguous()
        r2 = r.reshape(-1, D).contiguous() if r is not None else None
        rows = x2.shape[0]
        g = gamma if gamma.dtype == x2.dtype else gamma.to(x2.dtype)
        b = beta if beta.dtype == x2.dtype else beta.to(x2.dtype)
        y = torch.empty_like(x2)
        s = torch.empty_like(x2) if r2 is not None else x2
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        check(L.dfb_add_layernorm_fwd(ptr(x2), ptr(r2), ptr(g), ptr(b), ptr(s) if r2 is not None else None, ptr(y), ptr(mean),
                                      ptr(rstd), rows, D, eps, dt(x2), stream()), "dfb_add_layernorm_fwd")
        ctx.save_for_backward(s, g, mean, rstd)
        ctx.refs = (gamma, beta)
        ctx.has_r = r is not None
        ctx.shape = x.shape
        if r is not None:
            return s.view(x.shape), y.view(x.shape)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, *grads):
        L = _sigs()
        s, g, mean, rstd = ctx.saved_tensors
        gamma, beta = ctx.refs
        rows, D = s.shape
        if ctx.has_r:
            ds, dy = grads
        else:
            ds, dy = None, grads[0]
        if dy is None:
            dy = torch.zeros_like(s)
        dy2 = dy.reshape(rows, D).contiguous()
        ds2 = ds.reshape(rows, D).contiguous() if ds is not None else None
        dx = torch.empty_like(s)
        nblocks = int(L.dfb_layernorm_bwd_blocks(rows))
        # with a residual branch, dx is also the gradient of the Dense layer that produced the branch: its column sums are
        # that layer's bias gradient, produced here for free and handed over on the tensor (`ops.dense.take_colsum`)
        colsum = torch.empty(D, dtype=torch.float32, device=s.device) if (ctx.has_r and FUSE_BIAS_GRAD) else None
        part = torch.empty(nblocks * (3 if colsum is not None else 2) * D, dtype=torch.float32, device=s.device)
        sg, sb = getattr(gamma, "_dfb_grad_sink", None), getattr(beta, "_dfb_grad_sink", None)
        direct = sg is not None and sb is not None and sg.dtype == s.dtype and sb.dtype == s.dtype
        dgamma = sg if direct else torch.empty(D, dtype=s.dtype, device=s.device)
        dbeta = sb if direct else torch.empty(D, dtype=s.dtype, device=s.device)
        check(L.dfb_add_layernorm_bwd(ptr(dy2), ptr(ds2), ptr(s), ptr(g), ptr(mean), ptr(rstd), ptr(dx), ptr(dgamma), ptr(dbeta),
                                      ptr(colsum), ptr(part), rows, D, dt(s), stream()), "dfb_add_layernorm_bwd")
        if direct:
            gamma._dfb_grad_ready()
            beta._dfb_grad_ready()
            dgamma = dbeta = None
        else:
            dgamma, dbeta = dgamma.to(gamma.dtype), dbeta.to(beta.dtype)
        dxv = dx.view(ctx.shape)
        if not ctx.has_r:
            return dxv, None, dgamma, dbeta, None
        dr = dx.view(ctx.shape)                     # a second view object for the branch: it carries the column sums
        if colsum is not None:
            dr._dfb_colsum = colsum
        return dxv, dr, dgamma, dbeta, None


def layer_norm_cuda(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float) -> torch.Tensor:
    if not _supported(x):
        return F.layer_norm(x, (x.shape[-1],), gamma.to(x.dtype), beta.to(x.dtype), eps)
    return _AddLN.apply(x, None, gamma, beta, eps)


def add_layer_norm_cuda(x: torch.Tensor, r: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                        eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    if not _supported(x):
        s = x + r
        return s, F.layer_norm(s, (x.shape[-1],), gamma.to(x.dtype), beta.to(x.dtype), eps)
    return _AddLN.apply(x, r, gamma, beta, eps)


def permute_rows(src: torch.Tensor, dst: torch.Tensor, dims, src_strides, dst_strides, row_elems: int) -> None:
    """Copy rows of `row_elems` contiguous elements; strides are in elements."""
    L = _sigs()
    es = src.element_size()
    n = list(dims) + [1] * (4 - len(dims))
    ss = [int(v) * es for v in src_strides] + [0] * (4 - len(dims))
    ds = [int(v) * es for v in dst_strides] + [0] * (4 - len(dims))
    check(L.dfb_permute_rows(ptr(src), ptr(dst), *n, *ss, *ds, row_elems * es, stream()), "dfb_permute_rows")


# ------------------------------------------------------------------------------------------------ BatchNorm (channels last)
def _bn_supported(x: torch.Tensor, gamma: torch.Tensor) -> bool:
    vn = 8 if x.dtype == torch.bfloat16 else 4
    return x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and x.shape[-1] % vn == 0 and \
        gamma.dtype in (torch.bfloat16, torch.float32)


class _BatchNormFn(torch.autograd.Function):
    """y = act(BN(x) [+ residual]) over the last axis, batch statistics (training) — `csrc/batchnorm.cu`."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, momentum, eps, relu):
        L = _sigs()
        C = x.shape[-1]
        x2 = x.reshape(-1, C)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        r2 = None
        if residual is not None:
            r2 = residual.reshape(-1, C)
            if not r2.is_contiguous() or r2.dtype != x2.dtype:
                r2 = r2.to(x2.dtype).contiguous()
        rows = x2.shape[0]
        b = beta if beta.dtype == gamma.dtype else beta.to(gamma.dtype)
        y = torch.empty_like(x2)
        dev = x.device
        stats = torch.empty(2 * C, dtype=torch.float32, device=dev)           # saved mean | saved invstd
        gy = int(L.dfb_bn_partial_blocks(rows, C, dt(x2)))
        part = torch.empty(gy * 2 * C, dtype=torch.float32, device=dev)
        ab = torch.empty(2 * C, dtype=torch.float32, device=dev)
        direct = running_mean is not None and running_mean.dtype == torch.float32 and running_var.dtype == torch.float32 \
            and running_mean.is_cuda and running_mean.is_contiguous() and running_var.is_contiguous()
        check(L.dfb_bn_fwd_train(ptr(x2), ptr(r2), ptr(gamma), ptr(b), ptr(running_mean) if direct else None,
                                 ptr(running_var) if direct else None, ptr(y), ptr(stats), ptr(stats[C:]), ptr(part), ptr(ab), rows, C,
                                 eps, momentum, int(bool(relu)), dt(x2), dt(gamma), stream()), "dfb_bn_fwd_train")
        if running_mean is not None and not direct:                            # e.g. bf16 buffers of a `.to(bfloat16)` model
            n = float(rows)
            var = (1.0 / stats[C:].square() - eps).clamp_min_(0) * (n / max(n - 1.0, 1.0))
            running_mean.mul_(1 - momentum).add_(stats[:C].to(running_mean.dtype), alpha=momentum)
            running_var.mul_(1 - momentum).add_(var.to(running_var.dtype), alpha=momentum)
        ctx.save_for_backward(x2, y if relu else None, gamma, stats)
        ctx.refs = (gamma, beta)
        ctx.relu, ctx.has_res, ctx.shape = bool(relu), residual is not None, x.shape
        ctx.res_dtype = residual.dtype if residual is not None else None
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        L = _sigs()
        x2, y, g, stats = ctx.saved_tensors
        gamma, beta = ctx.refs
        rows, C = x2.shape
        dy2 = dy.reshape(rows, C)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dev = x2.device
        dx = torch.empty_like(x2)
        need_dres = ctx.has_res and ctx.relu
        dres = torch.empty_like(x2) if need_dres else None
        gy = int(L.dfb_bn_partial_blocks(rows, C, dt(x2)))
        part = torch.empty(gy * 2 * C, dtype=torch.float32, device=dev)
        coef = torch.empty(3 * C, dtype=torch.float32, device=dev)
        sg, sb = getattr(gamma, "_dfb_grad_sink", None), getattr(beta, "_dfb_grad_sink", None)
        direct = sg is not None and sb is not None and sg.dtype == g.dtype and sb.dtype == g.dtype
        dgamma = sg if direct else torch.empty(C, dtype=g.dtype, device=dev)
        dbeta = sb if direct else torch.empty(C, dtype=g.dtype, device=dev)
        check(L.dfb_bn_bwd(ptr(dy2), ptr(y), ptr(x2), ptr(g), ptr(stats), ptr(stats[C:]), ptr(dx), ptr(dres), ptr(dgamma), ptr(dbeta),
                           ptr(part), ptr(coef), rows, C, int(ctx.relu), dt(x2), dt(g), stream()), "dfb_bn_bwd")
        if direct:
            gamma._dfb_grad_ready()
            beta._dfb_grad_ready()
            dgamma = dbeta = None
        else:
            dgamma, dbeta = dgamma.to(gamma.dtype), dbeta.to(beta.dtype)
        d_res = None
        if ctx.has_res:
            d_res = (dres if need_dres else dy2).view(ctx.shape)
            if d_res.dtype != ctx.res_dtype:
                d_res = d_res.to(ctx.res_dtype)
        return dx.view(ctx.shape), d_res, dgamma, dbeta, None, None, None, None, None


def batch_norm_cuda(x, gamma, beta, running_mean, running_var, training: bool, momentum: float, eps: float,
                    activation=None, residual=None) -> torch.Tensor:
    """Channels-last BatchNorm + optional residual add + optional ReLU.  `momentum` is the weight of the NEW statistics."""
    relu = activation == "relu"
    if activation not in (None, "relu"):
        raise ValueError(f"batch_norm: unsupported activation {activation}")
    if training and _bn_supported(x, gamma):
        return _BatchNormFn.apply(x, residual, gamma, beta, running_mean, running_var, float(momentum), float(eps), relu)
    if (not training) and _bn_supported(x, gamma) and not torch.is_grad_enabled() and running_mean is not None \
            and running_mean.dtype == torch.float32 and running_var.dtype == torch.float32:
        L = _sigs()
        C = x.shape[-1]
        x2 = x.reshape(-1, C).contiguous()
        r2 = residual.reshape(-1, C).to(x2.dtype).contiguous() if residual is not None else None
        y = torch.empty_like(x2)
        ab = torch.empty(2 * C, dtype=torch.float32, device=x.device)
        b = beta if beta.dtype == gamma.dtype else beta.to(gamma.dtype)
        check(L.dfb_bn_fwd_eval(ptr(x2), ptr(r2), ptr(gamma), ptr(b), ptr(running_mean), ptr(running_var), ptr(y), ptr(ab),
                                x2.shape[0], C, eps, int(relu), dt(x2), dt(gamma), stream()), "dfb_bn_fwd_eval")
        return y.view(x.shape)
    return batch_norm_reference(x, gamma, beta, running_mean, running_var, training, momentum, eps, activation, residual)


def batch_norm_reference(x, gamma, beta, running_mean, running_var, training: bool, momentum: float, eps: float,
                         activation=None, residual=None) -> torch.Tensor:
    """Plain PyTorch statement of the same op (CPU path, eval-mode training graphs, and the baseline arm)."""
    if x.dim() == 4:                                  # NHWC storage viewed as NCHW channels_last: no copy
        y = F.batch_norm(x.permute(0, 3, 1, 2), running_mean, running_var, gamma, beta, training, momentum, eps).permute(0, 2, 3, 1)
    else:
        y = F.batch_norm(x, running_mean, running_var, gamma, beta, training, momentum, eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if activation == "relu" else y
