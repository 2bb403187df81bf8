"""Causal self-attention on the packed QKV projection.

Head dimension 64 in bf16 (GPT-2) runs the hand-written tcgen05 kernels of `csrc/attention_sm100.cu`: forward and
backward read `[B,T,3,H,D]` through 4-D TMA boxes and the backward writes `dqkv` in the same packed layout, so no head
split, merge or gradient pack exists any more (`_FlashAttention`).  Other shapes fall back to the library SDPA wrappers
below, kept from round 1: `_PackedAttention` hands the library strided views of
the packed projection output `[B,T,3,H,D]` (measured on B200: the cuDNN kernels take them at the speed of contiguous
operands and return the output in `[B,T,H,D]` memory order, so splitting and merging heads are views), and packs the
three gradients, which come back as separate `[B,T,H,D]` tensors, into `[B,T,3,H,D]` with the hand-written row-copy
kernel.  One copy per layer instead of eight.  `_SplitHeads` / `_MergeHeads` are the older explicit-permutation path,
kept for operands the packed path does not cover.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import c_int, c_void_p, check, ptr, stream


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


class _FlashAttention(torch.autograd.Function):
    """`qkv [B,T,3*H*64]` bf16 -> `[B,T,H*64]` through `dfb_attn_fwd` / `dfb_attn_bwd` (causal, scale 1/8)."""

    @staticmethod
    def forward(ctx, qkv, n_head):
        L = _sigs()
        B, T, C3 = qkv.shape
        D = C3 // (3 * n_head)
        q = qkv.detach()
        if not q.is_contiguous():
            q = q.contiguous()
        out = torch.empty((B, T, n_head * D), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((B, n_head, T), dtype=torch.float32, device=qkv.device)
        check(L.dfb_attn_fwd(ptr(q), ptr(out), ptr(lse), B, T, n_head, D, stream()), "dfb_attn_fwd")
        ctx.dims = (B, T, n_head, D)
        ctx.save_for_backward(q, out, lse)
        return out

    @staticmethod
    def backward(ctx, do):
        L = _sigs()
        q, out, lse = ctx.saved_tensors
        B, T, H, D = ctx.dims
        do = do.contiguous()
        if do.dtype != q.dtype:
            do = do.to(q.dtype)
        dqkv = torch.empty_like(q)
        scratch = _bwd_scratch(q.device, (B, T, H, D), int(L.dfb_attn_bwd_scratch_bytes(B, T, H, D)))
        from .norm import FUSE_BIAS_GRAD
        # the column sums of dqkv (= bias gradient of the QKV projection) come out of the same kernels: no pass over dqkv
        dbias = torch.empty(3 * H * D, dtype=torch.float32, device=q.device) if FUSE_BIAS_GRAD else None
        check(L.dfb_attn_bwd(ptr(q), ptr(out), ptr(do), ptr(lse), ptr(dqkv), ptr(dbias), ptr(scratch),
                             ctypes.c_size_t(scratch.numel()), B, T, H, D, stream()), "dfb_attn_bwd")
        if dbias is not None:
            dqkv._dfb_colsum = dbias
        return dqkv, None


_SCRATCH = {}


def _bwd_scratch(device: torch.device, shape, nbytes: int) -> torch.Tensor:
    """delta + the fp32 dQ accumulator of the backward kernel.  One zero-initialised buffer per device AND problem shape,
    shared by every layer of that shape (the kernel leaves the accumulator part at zero again, but where that part starts
    depends on the shape, so a buffer is never reused across shapes); created during the eager warm-up steps, i.e. before a
    CUDA graph capture of the step."""
    key = (device.type, device.index, tuple(shape))
    buf = _SCRATCH.get(key)
    if buf is None:
        if len(_SCRATCH) >= 8:                       # a handful of live shapes at most (training + evaluation)
            _SCRATCH.pop(next(iter(_SCRATCH)))
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _SCRATCH[key] = buf
    return buf


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_attn_sigs", False):
        L.dfb_attn_supported.restype = c_int
        L.dfb_attn_supported.argtypes = [c_int] * 4
        L.dfb_attn_fwd.restype = c_int
        L.dfb_attn_fwd.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]
        L.dfb_attn_bwd_scratch_bytes.restype = ctypes.c_size_t
        L.dfb_attn_bwd_scratch_bytes.argtypes = [c_int] * 4
        L.dfb_attn_bwd.restype = c_int
        L.dfb_attn_bwd.argtypes = [c_void_p] * 7 + [ctypes.c_size_t, c_int, c_int, c_int, c_int, c_void_p]
        L._dfb_attn_sigs = True
    return L


PACKED = True     # strided-view path (one gradient pack per layer); False: explicit head permutations
FLASH = os.environ.get("DFB_ATTN", "flash") == "flash"      # DFB_ATTN=sdpa: library attention for A/B runs


def causal_self_attention(qkv: torch.Tensor, n_head: int) -> torch.Tensor:
    """`qkv` `[B, T, 3*H*D]` (q | k | v along the last axis) -> `[B, T, H*D]`."""
    B, T, C3 = qkv.shape
    D = C3 // (3 * n_head)
    if qkv.is_cuda and FLASH and qkv.dtype == torch.bfloat16 and D == 64:
        return _FlashAttention.apply(qkv, n_head)
    if qkv.is_cuda and PACKED and (n_head * D * qkv.element_size()) % 16 == 0:
        return _PackedAttention.apply(qkv, n_head)
    if qkv.is_cuda and (D * qkv.element_size()) % 16 == 0:
        q, k, v = _SplitHeads.apply(qkv, n_head)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return _MergeHeads.apply(a)
    q, k, v = qkv.view(B, T, 3, n_head, D).permute(2, 0, 3, 1, 4)
    a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return a.transpose(1, 2).reshape(B, T, n_head * D)
