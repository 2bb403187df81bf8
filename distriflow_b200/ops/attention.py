"""Causal self-attention around the packed QKV projection.

The score/softmax/value product itself stays on `F.scaled_dot_product_attention` (attention is not one of the
hand-written hot ops named in BASELINE.json).  What this module owns is the data movement around it, which the step
profile showed to cost more than the attention forward itself.  `_PackedAttention` hands the library strided views of
the packed projection output `[B,T,3,H,D]` (measured on B200: the cuDNN kernels take them at the speed of contiguous
operands and return the output in `[B,T,H,D]` memory order, so splitting and merging heads are views), and packs the
three gradients, which come back as separate `[B,T,H,D]` tensors, into `[B,T,3,H,D]` with the hand-written row-copy
kernel.  One copy per layer instead of eight.  `_SplitHeads` / `_MergeHeads` are the older explicit-permutation path,
kept for operands the packed path does not cover.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def _bthd_dense(t: torch.Tensor, B: int, H: int, T: int, D: int) -> bool:
    return tuple(t.shape) == (B, H, T, D) and t.stride() == (T * H * D, D, H * D, 1)


class _PackedAttention(torch.autograd.Function):
    """`qkv [B,T,3*H*D]` -> `[B,T,H*D]`; backward returns the packed `dqkv`."""

    @staticmethod
    def forward(ctx, qkv, n_head):
        B, T, C3 = qkv.shape
        D = C3 // (3 * n_head)
        base = qkv.detach().contiguous().view(B, T, 3, n_head, D)
        q, k, v = (base[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        ctx.dims = (B, T, n_head, D)
        if ctx.needs_input_grad[0]:
            q, k, v = (t.requires_grad_() for t in (q, k, v))
            with torch.enable_grad():
                o = F.scaled_dot_product_attention(q, k, v, is_causal=True)      # [B,H,T,D] logical
            ctx.inner = (q, k, v, o)
        else:                                                                    # inference: no graph, nothing retained
            o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
            ctx.inner = None
        od = o.detach()
        if _bthd_dense(od, B, n_head, T, D):
            return od.permute(0, 2, 1, 3).reshape(B, T, n_head * D)              # a view: memory is already [B,T,H,D]
        return _MergeHeads.apply(od)

    @staticmethod
    def backward(ctx, do):
        from .norm import permute_rows
        B, T, H, D = ctx.dims
        q, k, v, o = ctx.inner
        ctx.inner = None
        g = do.contiguous().view(B, T, H, D).permute(0, 2, 1, 3)
        if not _bthd_dense(o, B, H, T, D):
            g = g.contiguous()
        dq, dk, dv = torch.autograd.grad(o, (q, k, v), g)
        HD, C3 = H * D, 3 * H * D
        dqkv = torch.empty((B, T, C3), dtype=do.dtype, device=do.device)
        rb = 8 if T % 8 == 0 else 1
        for w, gw in enumerate((dq, dk, dv)):
            if not _bthd_dense(gw, B, H, T, D):                                  # library returned another layout: normalise
                gw = gw.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
            permute_rows(gw, dqkv[:, :, w * HD:], (B, T // rb, rb), (T * HD, rb * HD, HD), (T * C3, rb * C3, C3), HD)
        return dqkv, None


class _SplitHeads(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, n_head):
        from .norm import permute_rows
        B, T, C3 = qkv.shape
        D = C3 // (3 * n_head)
        qkv = qkv.contiguous()
        out = torch.empty((3, B, n_head, T, D), dtype=qkv.dtype, device=qkv.device)
        for w in range(3):      # rows indexed (b, h, t): src [B,T,3,H,D] -> dst [3,B,H,T,D]
            permute_rows(qkv[:, :, w * n_head * D:], out[w], (B, n_head, T), (T * C3, D, C3), (n_head * T * D, T * D, D), D)
        ctx.dims = (B, T, n_head, D)
        return out[0], out[1], out[2]

    @staticmethod
    def backward(ctx, dq, dk, dv):
        from .norm import permute_rows
        B, T, H, D = ctx.dims
        C3 = 3 * H * D
        dqkv = torch.empty((B, T, C3), dtype=dq.dtype, device=dq.device)
        for w, g in enumerate((dq, dk, dv)):
            g = g.contiguous()
            permute_rows(g, dqkv[:, :, w * H * D:], (B, H, T), (H * T * D, T * D, D), (T * C3, D, C3), D)
        return dqkv, None


class _MergeHeads(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a):                       # [B,H,T,D] -> [B,T,H*D]
        from .norm import permute_rows
        B, H, T, D = a.shape
        a = a.contiguous()
        out = torch.empty((B, T, H * D), dtype=a.dtype, device=a.device)
        permute_rows(a, out, (B, H, T), (H * T * D, T * D, D), (T * H * D, D, H * D), D)
        ctx.dims = (B, H, T, D)
        return out

    @staticmethod
    def backward(ctx, do):                     # [B,T,H*D] -> [B,H,T,D]
        from .norm import permute_rows
        B, H, T, D = ctx.dims
        do = do.contiguous()
        out = torch.empty((B, H, T, D), dtype=do.dtype, device=do.device)
        permute_rows(do, out, (B, H, T), (T * H * D, D, H * D), (H * T * D, T * D, D), D)
        return out


PACKED = True     # strided-view path (one gradient pack per layer); False: explicit head permutations


def causal_self_attention(qkv: torch.Tensor, n_head: int) -> torch.Tensor:
    """`qkv` `[B, T, 3*H*D]` (q | k | v along the last axis) -> `[B, T, H*D]`."""
    B, T, C3 = qkv.shape
    D = C3 // (3 * n_head)
    if qkv.is_cuda and PACKED and (n_head * D * qkv.element_size()) % 16 == 0:
        return _PackedAttention.apply(qkv, n_head)
    if qkv.is_cuda and (D * qkv.element_size()) % 16 == 0:
        q, k, v = _SplitHeads.apply(qkv, n_head)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return _MergeHeads.apply(a)
    q, k, v = qkv.view(B, T, 3, n_head, D).permute(2, 0, 3, 1, 4)
    a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return a.transpose(1, 2).reshape(B, T, n_head * D)
