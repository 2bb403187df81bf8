"""NHWC Conv2D / MaxPool on CUDA through the hand-written implicit-GEMM kernels.

Filters are HWIO and activations NHWC, as tf.js stores them, so checkpoints of
`/root/reference/experiment/mnist/model.json` map without a transpose.  bf16 with channel counts that are
multiples of 16 takes the tcgen05 implicit-GEMM kernel (`csrc/conv_sm100.cu`, TMA gathers the shifted NHWC
slab of every filter tap); the rest takes the CUDA-core implicit GEMM (`csrc/conv_simt.cu`).
Replaces K2/K4 of SURVEY §2.5.
"""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch

from . import _lib
from ._lib import c_int, c_void_p, check, dt, ptr, stream
from .dense import ACT, colsum


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_conv_sigs", False):
        L.dfb_conv2d_simt.restype = c_int
        L.dfb_conv2d_simt.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 18 + [c_void_p]
        L.dfb_maxpool_nhwc.restype = c_int
        L.dfb_maxpool_nhwc.argtypes = [c_void_p, c_void_p] + [c_int] * 13 + [c_void_p]
        L.dfb_maxpool_nhwc_bwd.restype = c_int
        L.dfb_maxpool_nhwc_bwd.argtypes = [c_void_p, c_void_p, c_void_p] + [c_int] * 13 + [c_void_p]
        L.dfb_conv2d_tc.restype = c_int
        L.dfb_conv2d_tc.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 18 + [c_void_p, ctypes.c_size_t, c_void_p]
        L.dfb_col2im_nhwc.restype = c_int
        L.dfb_col2im_nhwc.argtypes = [c_void_p, c_void_p] + [c_int] * 13 + [c_void_p]
        L.dfb_im2col_nhwc.restype = c_int
        L.dfb_im2col_nhwc.argtypes = [c_void_p, c_void_p] + [c_int] * 13 + [c_void_p]
        L.dfb_conv2d_tc_supported.restype = c_int
        L.dfb_conv2d_tc_supported.argtypes = [c_int] * 12
        L._dfb_conv_sigs = True
    return L


def _geometry(H: int, W: int, kh: int, kw: int, strides: Tuple[int, int], padding: str):
    from . import same_padding
    sh, sw = strides
    if padding == "same":
        pt, _ = same_padding(H, kh, sh)
        pl, _ = same_padding(W, kw, sw)
        OH, OW = -(-H // sh), -(-W // sw)
    elif padding == "valid":
        pt = pl = 0
        OH, OW = (H - kh) // sh + 1, (W - kw) // sw + 1
    else:
        raise ValueError(f"padding {padding}")
    return OH, OW, pt, pl


def _tc_eligible(mode: int, x: torch.Tensor, geom) -> bool:
    if x.dtype != torch.bfloat16 or FORCE_SIMT:
        return False
    (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = geom
    return bool(_sigs().dfb_conv2d_tc_supported(mode, N, H, W, C, OH, OW, K, kh, kw, sh, sw))


FORCE_SIMT = False      # tests flip this to compare the two implementations


def _conv_call(mode: int, a, b, out, bias, geom, act, ksplit=1):
    L = _sigs()
    (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = geom
    check(L.dfb_conv2d_simt(mode, ptr(a), ptr(b), ptr(out), ptr(bias), dt(bias) if bias is not None else 0,
                            N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl, dt(a), 1 if out.dtype == torch.float32 else 0,
                            ACT[act], ksplit, stream()), "dfb_conv2d_simt")
    return out


def _conv_tc(mode: int, a, b, out, bias, geom, act):
    L = _sigs()
    (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = geom
    check(L.dfb_conv2d_tc(mode, ptr(a), ptr(b), ptr(out), ptr(bias), dt(bias) if bias is not None else 0,
                          N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl, dt(a), 1 if out.dtype == torch.float32 else 0,
                          ACT[act], 0, None, ctypes.c_size_t(0), stream()), "dfb_conv2d_tc")
    return out


def _im2col(x, geom, kpad):
    L = _sigs()
    (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = geom
    out = torch.empty((N * OH * OW, kpad), device=x.device, dtype=torch.bfloat16)
    check(L.dfb_im2col_nhwc(ptr(x), ptr(out), N, H, W, C, OH, OW, kh, kw, sh, sw, pt, pl, kpad, stream()), "dfb_im2col_nhwc")
    return out


class _ThinConvFn(torch.autograd.Function):
    """Convolution with very few input channels (ResNet stems, C_in = 3): bf16 patch matrix + the tcgen05 Dense GEMM for the
    forward and the filter gradient; the patch matrix is rebuilt in backward instead of being kept."""

    @staticmethod
    def forward(ctx, x, kernel, bias, strides, padding, activation):
        from .dense import _init_sigs, gemm
        _init_sigs()
        x = x.contiguous()
        N, H, W, C = x.shape
        kh, kw, _, K = kernel.shape
        OH, OW, pt, pl = _geometry(H, W, kh, kw, strides, padding)
        geom = (N, H, W, C, OH, OW, K, kh, kw, strides[0], strides[1], pt, pl)
        kreal = kh * kw * C
        kpad = -(-kreal // 64) * 64
        wmat = torch.zeros((kpad, K), device=x.device, dtype=torch.bfloat16)
        wmat[:kreal] = kernel.reshape(kreal, K).to(torch.bfloat16)
        patches = _im2col(x, geom, kpad)
        y = gemm(patches, wmat, a_mn=False, b_mn=True, M=N * OH * OW, N=K, K=kpad, bias=bias, act=activation)
        ctx.geom, ctx.activation, ctx.kpad = geom, activation, kpad
        ctx.kernel_ref, ctx.bias_ref = kernel, bias
        ctx.save_for_backward(x, wmat, y if activation == "relu" else None)
        return y.view(N, OH, OW, K)

    @staticmethod
    def backward(ctx, dy):
        from .dense import act_op, gemm
        x, wmat, y = ctx.saved_tensors
        (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = ctx.geom
        kreal = kh * kw * C
        dy2 = dy.reshape(-1, K).contiguous()
        if ctx.activation == "relu":
            dy2 = act_op(dy2, y, 2)
        kernel, bias = ctx.kernel_ref, ctx.bias_ref
        dx = dw = db = None
        # data gradient first: it reads the live filter, and `_dfb_grad_ready()` below may start the bucket kernel that overwrites it
        if ctx.needs_input_grad[0]:
            if sh == 1 and sw == 1 and K % 8 == 0 and C % 8 == 0:
                # data gradient = stride-1 correlation of dY with the flipped, transposed filter: same patch-matrix GEMM
                pth, plw = kh - 1 - pt, kw - 1 - pl
                g2 = (N, OH, OW, K, H, W, C, kh, kw, 1, 1, pth, plw)
                kreal2 = kh * kw * K
                kpad2 = -(-kreal2 // 64) * 64
                wflip = torch.zeros((kpad2, C), device=x.device, dtype=torch.bfloat16)
                wflip[:kreal2] = kernel.detach().to(torch.bfloat16).flip(0, 1).permute(0, 1, 3, 2).reshape(kreal2, C)
                pdy = _im2col(dy2.view(N, OH, OW, K), g2, kpad2)
                dx = gemm(pdy, wflip, a_mn=False, b_mn=True, M=N * H * W, N=C, K=kpad2).view(N, H, W, C)
            elif C % 8 == 0 and K % 8 == 0:
                # strided: patch gradient dP = dY @ Wmat^T on the tensor cores, then the gather form of col2im
                dpatch = gemm(dy2, wmat, a_mn=False, b_mn=False, M=N * OH * OW, N=ctx.kpad, K=K)
                dx = torch.empty_like(x)
                check(_sigs().dfb_col2im_nhwc(ptr(dpatch), ptr(dx), N, H, W, C, OH, OW, kh, kw, sh, sw, pt, pl, ctx.kpad, stream()),
                      "dfb_col2im_nhwc")
            else:
                dx = torch.empty_like(x)
                _conv_call(1, dy2.view(N, OH, OW, K), ctx.kernel_ref.to(x.dtype).contiguous(), dx, None, ctx.geom, None)
        if ctx.needs_input_grad[1]:
            patches = _im2col(x, ctx.geom, ctx.kpad)
            dwm = gemm(patches, dy2, a_mn=True, b_mn=True, M=ctx.kpad, N=K, K=N * OH * OW, out_dtype=torch.float32)
            dw_full = dwm[:kreal].reshape(kh, kw, C, K)
            sink = getattr(kernel, "_dfb_grad_sink", None)
            if sink is not None:
                sink.view(kh, kw, C, K).copy_(dw_full)
                kernel._dfb_grad_ready()
            else:
                dw = dw_full.to(kernel.dtype)
        if bias is not None and ctx.needs_input_grad[2]:
            db = colsum(dy2)
            sink = getattr(bias, "_dfb_grad_sink", None)
            if sink is not None:
                sink.copy_(db)
                bias._dfb_grad_ready()
                db = None
            else:
                db = db.to(bias.dtype)
        return dx, dw, db, None, None, None


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, bias, strides, padding, activation):
        x = x.contiguous()
        N, H, W, C = x.shape
        kh, kw, C2, K = kernel.shape
        assert C == C2, f"input has {C} channels, filter expects {C2}"
        OH, OW, pt, pl = _geometry(H, W, kh, kw, strides, padding)
        geom = (N, H, W, C, OH, OW, K, kh, kw, strides[0], strides[1], pt, pl)
        w = kernel if kernel.dtype == x.dtype else kernel.to(x.dtype)
        w = w.contiguous()
        y = torch.empty((N, OH, OW, K), device=x.device, dtype=x.dtype)
        if _tc_eligible(0, x, geom):
            _conv_tc(0, x, w, y, bias, geom, activation)
        else:
            _conv_call(0, x, w, y, bias, geom, activation)
        ctx.geom, ctx.activation = geom, activation
        ctx.kernel_ref, ctx.bias_ref = kernel, bias
        ctx.save_for_backward(x, w, y if activation == "relu" else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .dense import act_op
        x, w, y = ctx.saved_tensors
        geom = ctx.geom
        (N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl) = geom
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        if ctx.activation == "relu":
            dy = act_op(dy, y, 2)
        dx = dw = db = None
        kernel, bias = ctx.kernel_ref, ctx.bias_ref
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            if _tc_eligible(1, x, geom):
                _conv_tc(1, dy, w, dx, None, geom, None)
            else:
                _conv_call(1, dy, w, dx, None, geom, None)
        if ctx.needs_input_grad[1]:
            pixels = N * OH * OW
            tiles = (-(-(kh * kw * C) // 64)) * (-(-K // 64))
            ksplit = max(1, min(64, (148 * 4) // max(tiles, 1), pixels // 256))
            dw32 = torch.empty((kh, kw, C, K), device=x.device, dtype=torch.float32)
            if _tc_eligible(2, x, geom):
                _conv_tc(2, x, dy, dw32, None, geom, None)
            else:
                _conv_call(2, x, dy, dw32, None, geom, None, ksplit)
            sink = getattr(kernel, "_dfb_grad_sink", None)
            if sink is not None:
                sink.view(kh, kw, C, K).copy_(dw32)
                kernel._dfb_grad_ready()
            else:
                dw = dw32.to(kernel.dtype)
        if bias is not None and ctx.needs_input_grad[2]:
            db = colsum(dy.reshape(-1, K))
            sink = getattr(bias, "_dfb_grad_sink", None)
            if sink is not None:
                sink.copy_(db)
                bias._dfb_grad_ready()
                db = None
            else:
                db = db.to(bias.dtype)
        return dx, dw, db, None, None, None


def conv2d_cuda(x, kernel, bias, strides, padding, activation):
    C, K = kernel.shape[2], kernel.shape[3]
    thin = C < 64 or K < 64            # no 128-byte channel rows for the TMA implicit-GEMM path
    if x.dtype == torch.bfloat16 and thin and K % 16 == 0 and not FORCE_SIMT and x.shape[0] * x.shape[1] * x.shape[2] >= 4096:
        return _ThinConvFn.apply(x, kernel, bias, tuple(strides), padding, activation)
    return _Conv2dFn.apply(x, kernel, bias, tuple(strides), padding, activation)


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, pool, strides, padding):
        L = _sigs()
        x = x.contiguous()
        N, H, W, C = x.shape
        OH, OW, pt, pl = _geometry(H, W, pool[0], pool[1], strides, padding)
        y = torch.empty((N, OH, OW, C), device=x.device, dtype=x.dtype)
        check(L.dfb_maxpool_nhwc(ptr(x), ptr(y), N, H, W, C, OH, OW, pool[0], pool[1], strides[0], strides[1], pt, pl, dt(x),
                                 stream()), "dfb_maxpool_nhwc")
        ctx.args = (N, H, W, C, OH, OW, pool[0], pool[1], strides[0], strides[1], pt, pl)
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _sigs()
        (x,) = ctx.saved_tensors
        N, H, W, C, OH, OW, kh, kw, sh, sw, pt, pl = ctx.args
        dy = dy.contiguous()
        overlap = sh < kh or sw < kw
        if overlap and x.dtype != torch.float32:
            x32, dy32 = x.float(), dy.float()
            dx = torch.empty_like(x32)
            check(L.dfb_maxpool_nhwc_bwd(ptr(x32), ptr(dy32), ptr(dx), *ctx.args, 0, stream()), "dfb_maxpool_nhwc_bwd")
            return dx.to(x.dtype), None, None, None
        dx = torch.empty_like(x)
        check(L.dfb_maxpool_nhwc_bwd(ptr(x), ptr(dy), ptr(dx), *ctx.args, dt(x), stream()), "dfb_maxpool_nhwc_bwd")
        return dx, None, None, None


def maxpool2d_cuda(x, pool, strides, padding):
    return _MaxPoolFn.apply(x, tuple(pool), tuple(strides), padding)
