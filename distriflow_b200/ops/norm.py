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
        L.dfb_add_layernorm_bwd.restype = c_int
        L.dfb_add_layernorm_bwd.argtypes = [c_void_p] * 10 + [c_int, c_int, c_int, c_void_p]
        L.dfb_permute_rows.restype = c_int
        L.dfb_permute_rows.argtypes = [c_void_p, c_void_p] + [c_int] * 4 + [ctypes.c_long] * 8 + [c_int, c_void_p]
        L._dfb_norm_sigs = True
    return L


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
        part = torch.empty(nblocks * 2 * D, dtype=torch.float32, device=s.device)
        sg, sb = getattr(gamma, "_dfb_grad_sink", None), getattr(beta, "_dfb_grad_sink", None)
        direct = sg is not None and sb is not None and sg.dtype == s.dtype and sb.dtype == s.dtype
        dgamma = sg if direct else torch.empty(D, dtype=s.dtype, device=s.device)
        dbeta = sb if direct else torch.empty(D, dtype=s.dtype, device=s.device)
        check(L.dfb_add_layernorm_bwd(ptr(dy2), ptr(ds2), ptr(s), ptr(g), ptr(mean), ptr(rstd), ptr(dx), ptr(dgamma), ptr(dbeta),
                                      ptr(part), rows, D, dt(s), stream()), "dfb_add_layernorm_bwd")
        if direct:
            gamma._dfb_grad_ready()
            beta._dfb_grad_ready()
            dgamma = dbeta = None
        else:
            dgamma, dbeta = dgamma.to(gamma.dtype), dbeta.to(beta.dtype)
        dxv = dx.view(ctx.shape)
        return dxv, (dxv if ctx.has_r else None), dgamma, dbeta, None


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
