"""Causal self-attention around the packed QKV projection.

The score/softmax/value product itself stays on `F.scaled_dot_product_attention` (attention is not one of the
hand-written hot ops named in BASELINE.json).  What this module owns is the data movement around it, which the step
profile showed to cost more than the attention kernels: the packed projection output `[B,T,3,H,D]` is split into
contiguous `[B,H,T,D]` q/k/v by one hand-written permutation kernel (backward: one pack of dq/dk/dv into `[B,T,3,H,D]`,
replacing torch's `cat`), and the attention output is merged back to `[B,T,H*D]` by the same kernel (replacing the
generic strided copy).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


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


def causal_self_attention(qkv: torch.Tensor, n_head: int) -> torch.Tensor:
    """`qkv` `[B, T, 3*H*D]` (q | k | v along the last axis) -> `[B, T, H*D]`."""
    B, T, C3 = qkv.shape
    D = C3 // (3 * n_head)
    if qkv.is_cuda and (D * qkv.element_size()) % 16 == 0:
        q, k, v = _SplitHeads.apply(qkv, n_head)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return _MergeHeads.apply(a)
    q, k, v = qkv.view(B, T, 3, n_head, D).permute(2, 0, 3, 1, 4)
    a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return a.transpose(1, 2).reshape(B, T, n_head * D)
