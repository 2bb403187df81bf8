"""Dense layer on CUDA: forward, dgrad and wgrad through the hand-written GEMMs.

`weight` is `[in, out]` (tf.js Dense kernel layout), so no transpose is materialised:
the three products differ only in which operand is MN-major (see `csrc/gemm_sm100.cu`).
bf16 with 16-byte-aligned pitches runs the tcgen05/TMEM/TMA kernel; everything else (fp32 MNIST models,
N = 10 / N = 5 heads) runs the CUDA-core kernel.  Replaces K1/K4 of SURVEY §2.5
(`/root/reference/src/common/models.ts:137-142`).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import c_int, c_size_t, c_void_p, check, dt, ptr, stream

ACT = {None: 0, "relu": 1, "gelu": 2}
NUM_SMS = 148
_MAX_CTAS = 0            # 0 = all SMs; the engine lowers this while a comm kernel needs SMs


def set_max_ctas(n: int) -> None:
    global _MAX_CTAS
    _MAX_CTAS = int(n)


def _tc_ok(*mats: torch.Tensor) -> bool:
    for t in mats:
        if t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1 or t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
            return False
    return True


TWO_CTA = True           # use the cta_group::2 kernel (256x256 tiles on SM pairs) where it fills the machine


def pick_tile(M: int, N: int, K: int) -> Tuple[int, int]:
    """(block_n, splits).  block_n = 512 selects the 2-CTA kernel (256 x 256 tile per SM pair)."""
    pairs = NUM_SMS // 2
    kb = -(-K // 64)
    if TWO_CTA and M >= 256 and N >= 256 and _MAX_CTAS == 0:
        t2 = -(-M // 256) * -(-N // 256)
        if t2 >= pairs:
            return 512, 1
        if kb >= 32 and N % 4 == 0:          # short MN, long K (weight gradients): split K over the pairs
            sp = max(1, min(pairs // t2, kb // 8, 16))
            if t2 * sp * 10 >= pairs * 6:
                return 512, sp
    mb = -(-M // 128)
    t256 = mb * -(-N // 256)
    bn = 256 if (t256 >= NUM_SMS and N >= 256) else 128
    splits = _lib.lib().dfb_gemm_pick_splits(M, N, K, bn, NUM_SMS)
    return bn, int(splits)


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool, b_mn: bool, M: int, N: int, K: int,
         bias: Optional[torch.Tensor] = None, act: Optional[str] = None, out: Optional[torch.Tensor] = None,
         out_dtype: Optional[torch.dtype] = None, block_n: Optional[int] = None, splits: Optional[int] = None) -> torch.Tensor:
    """D[M,N] = act(A·Bᵀ + bias).  A is `[M,K]` memory if not a_mn else `[K,M]`; B is `[N,K]` if not b_mn else `[K,N]`."""
    L = _lib.lib()
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype or a.dtype)
    assert out.stride(1) == 1
    use_tc = _tc_ok(a, b) and out.dtype in (torch.bfloat16, torch.float32) and \
        (out.stride(0) * out.element_size()) % 16 == 0 and out.data_ptr() % 16 == 0 and N >= 16
    if bias is not None and bias.dtype not in (torch.float32, torch.bfloat16):
        bias = bias.float()
    if use_tc:
        bn, sp = pick_tile(M, N, K)
        bn = block_n or bn
        sp = sp if splits is None else splits
        if N % 4 != 0:
            sp = 1
        ws, ws_bytes = None, 0
        if sp > 1:
            ws_bytes = int(L.dfb_gemm_workspace_bytes(M, N, sp))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=a.device)
        check(L.dfb_gemm_bf16(ptr(a), ptr(b), ptr(out), ptr(bias), c_int(dt(bias) if bias is not None else 0),
                              M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), dt(out), ACT[act],
                              bn, sp, ptr(ws), c_size_t(ws_bytes), _MAX_CTAS, stream()), "dfb_gemm_bf16")
        return out
    # CUDA-core path: general strides
    if a.dtype != b.dtype:
        b = b.to(a.dtype)
    sa_m, sa_k = (a.stride(1), a.stride(0)) if a_mn else (a.stride(0), a.stride(1))
    sb_n, sb_k = (b.stride(1), b.stride(0)) if b_mn else (b.stride(0), b.stride(1))
    check(L.dfb_gemm_simt(ptr(a), ptr(b), ptr(out), ptr(bias), c_int(dt(bias) if bias is not None else 0), M, N, K,
                          ctypes.c_long(sa_m), ctypes.c_long(sa_k), ctypes.c_long(sb_n), ctypes.c_long(sb_k),
                          ctypes.c_long(out.stride(0)), dt(a), dt(b), dt(out), ACT[act], 1, stream()), "dfb_gemm_simt")
    return out


def ag_gemm(x2: torch.Tensor, w: torch.Tensor, plan, bias: Optional[torch.Tensor] = None, act: Optional[str] = None) -> torch.Tensor:
    """Forward GEMM `act(x2 @ w + bias)` that also gathers the remote shards of `w` (see parallel/ag_fusion.py)."""
    L = _lib.lib()
    M, K = x2.shape
    N = w.shape[1]
    out = torch.empty((M, N), device=x2.device, dtype=x2.dtype)
    bn = 256 if (-(-M // 128) * -(-N // 256) >= NUM_SMS and N >= 256) else 128
    if bias is not None and bias.dtype not in (torch.float32, torch.bfloat16):
        bias = bias.float()
    check(L.dfb_ag_gemm_bf16(ptr(x2), ptr(w), ptr(out), ptr(bias), c_int(dt(bias) if bias is not None else 0), M, N, K,
                             x2.stride(0), w.stride(0), out.stride(0), dt(out), ACT[act], bn, ptr(plan.pieces), ptr(plan.chunk0),
                             plan.num_pieces, ptr(plan.expected), ptr(plan.ready), plan.rot_kb, stream()), "dfb_ag_gemm_bf16")
    return out


def gemm_aux(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool, b_mn: bool, M: int, N: int, K: int, aux: torch.Tensor, aux_mode: int,
             bias: Optional[torch.Tensor] = None, act: Optional[str] = None) -> torch.Tensor:
    """tcgen05 GEMM whose epilogue also writes the pre-activation to `aux` (mode 1) or multiplies by gelu'(aux) (mode 2)."""
    L = _lib.lib()
    out = torch.empty((M, N), device=a.device, dtype=torch.bfloat16)
    bn, sp = pick_tile(M, N, K)
    if sp != 1:
        bn = 256 if (-(-M // 128) * -(-N // 256) >= NUM_SMS and N >= 256) else 128
    if bias is not None and bias.dtype not in (torch.float32, torch.bfloat16):
        bias = bias.float()
    check(L.dfb_gemm_bf16_aux(ptr(a), ptr(b), ptr(out), ptr(bias), c_int(dt(bias) if bias is not None else 0), M, N, K,
                              a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), ACT[act], bn, ptr(aux),
                              ctypes.c_long(aux.stride(0)), aux_mode, _MAX_CTAS, stream()), "dfb_gemm_bf16_aux")
    return out


def gemm_dgelu_colsum(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool, b_mn: bool, M: int, N: int, K: int, aux: torch.Tensor):
    """`(a·bᵀ) * gelu'(aux)` and, out of the same epilogue, its column sums (fp32 `[N]`): the pre-activation gradient of an
    MLP and the bias gradient of its first layer in one kernel.  Returns (out, None) where the 2-CTA kernel does not apply."""
    bn, sp = pick_tile(M, N, K)
    if bn != 512 or sp != 1 or not FUSE_DGELU_COLSUM:
        return gemm_aux(a, b, a_mn=a_mn, b_mn=b_mn, M=M, N=N, K=K, aux=aux, aux_mode=2), None
    L = _lib.lib()
    out = torch.empty((M, N), device=a.device, dtype=torch.bfloat16)
    cs = torch.empty(N, dtype=torch.float32, device=a.device)
    nbytes = int(L.dfb_gemm_colsum_part_bytes(M, N))
    part = torch.empty(nbytes // 4, dtype=torch.float32, device=a.device)
    check(L.dfb_gemm_bf16_dgelu_colsum(ptr(a), ptr(b), ptr(out), M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn),
                                       int(b_mn), ptr(aux), ctypes.c_long(aux.stride(0)), ptr(cs), ptr(part), c_size_t(nbytes),
                                       _MAX_CTAS, stream()), "dfb_gemm_bf16_dgelu_colsum")
    return out, cs


def take_colsum(dy: torch.Tensor, N: int) -> Optional[torch.Tensor]:
    """Column sums of `dy` that its producer already computed (fp32 `[N]`), or None.  The producer (LayerNorm backward for
    the residual branch, attention backward for the packed QKV gradient) hangs them on the gradient tensor it returns."""
    cs = getattr(dy, "_dfb_colsum", None)
    if cs is not None and cs.numel() == N and cs.device == dy.device:
        return cs
    return None


def colsum(x2d: torch.Tensor) -> torch.Tensor:
    L = _lib.lib()
    M, N = x2d.shape
    out = torch.empty(N, dtype=torch.float32, device=x2d.device)
    vn = 8 if x2d.dtype == torch.bfloat16 else 4
    if x2d.dtype not in (torch.bfloat16, torch.float32) or x2d.stride(0) % vn != 0 or x2d.stride(1) != 1 \
            or x2d.data_ptr() % 16 != 0:
        return x2d.float().sum(0)
    check(L.dfb_colsum(ptr(x2d), ptr(out), M, N, ctypes.c_long(x2d.stride(0)), dt(x2d), stream()), "dfb_colsum")
    return out


def act_op(a: torch.Tensor, b: Optional[torch.Tensor], mode: int) -> torch.Tensor:
    L = _lib.lib()
    a = a.contiguous()
    b = b.contiguous() if b is not None else a
    out = torch.empty_like(a)
    check(L.dfb_act(ptr(a), ptr(b), ptr(out), ctypes.c_long(a.numel()), dt(a), mode, stream()), "dfb_act")
    return out


def _init_sigs() -> None:
    L = _lib.lib()
    if getattr(L, "_dfb_linear_sigs", False):
        return
    L.dfb_gemm_workspace_bytes.restype = ctypes.c_size_t
    L.dfb_gemm_workspace_bytes.argtypes = [c_int, c_int, c_int]
    L.dfb_gemm_pick_splits.restype = c_int
    L.dfb_gemm_pick_splits.argtypes = [c_int] * 5
    L.dfb_gemm_bf16.restype = c_int
    L.dfb_gemm_bf16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_int, c_void_p]
    L.dfb_gemm_colsum_part_bytes.restype = ctypes.c_size_t
    L.dfb_gemm_colsum_part_bytes.argtypes = [c_int, c_int]
    L.dfb_gemm_bf16_dgelu_colsum.restype = c_int
    L.dfb_gemm_bf16_dgelu_colsum.argtypes = [c_void_p, c_void_p, c_void_p] + [c_int] * 8 + [c_void_p, ctypes.c_long, c_void_p, c_void_p,
                                                                                   c_size_t, c_int, c_void_p]
    L.dfb_gemm_simt.restype = c_int
    L.dfb_gemm_simt.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int] + [ctypes.c_long] * 5 + \
        [c_int] * 5 + [c_void_p]
    L.dfb_ag_gemm_bf16.restype = c_int
    L.dfb_ag_gemm_bf16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 10 + [c_void_p, c_void_p, c_int, c_void_p,
                                                                                          c_void_p, c_int, c_void_p]
    L.dfb_gemm_bf16_aux.restype = c_int
    L.dfb_gemm_bf16_aux.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 11 + [c_void_p, ctypes.c_long, c_int, c_int,
                                                                                           c_void_p]
    L.dfb_colsum.restype = c_int
    L.dfb_colsum.argtypes = [c_void_p, c_void_p, c_int, c_int, ctypes.c_long, c_int, c_void_p]
    L.dfb_act.restype = c_int
    L.dfb_act.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_long, c_int, c_int, c_void_p]
    L._dfb_linear_sigs = True


class _LinearFn(torch.autograd.Function):
    """`transposed=False`: weight `[in, out]` (tf.js Dense).  `transposed=True`: weight `[out, in]` (tied LM head)."""

    @staticmethod
    def forward(ctx, x, weight, bias, activation, transposed):
        _init_sigs()
        if transposed:
            N, K = weight.shape
        else:
            K, N = weight.shape
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        w = weight if weight.dtype == x2.dtype else weight.to(x2.dtype)
        M = x2.shape[0]
        saved_act = None
        plan = getattr(weight, "_dfb_ag", None)
        if plan is not None and plan.num_pieces > 0 and not transposed and w is weight and _tc_ok(x2, w) and N >= 16:
            # the all-gather of this kernel is fused into its first consumer: this GEMM pulls the remote shards itself
            if activation == "gelu":
                pre = ag_gemm(x2, w, plan, bias)
                y = act_op(pre, None, 0)
                saved_act = pre
            else:
                y = ag_gemm(x2, w, plan, bias, activation)
                if activation == "relu":
                    saved_act = y
        elif activation == "gelu":
            pre = gemm(x2, w, a_mn=False, b_mn=not transposed, M=M, N=N, K=K, bias=bias)
            y = act_op(pre, None, 0)
            saved_act = pre
        else:
            y = gemm(x2, w, a_mn=False, b_mn=not transposed, M=M, N=N, K=K, bias=bias, act=activation)
            if activation == "relu":
                saved_act = y
        ctx.activation = activation
        ctx.has_bias = bias is not None
        ctx.bias_ref = bias
        ctx.weight_ref = weight
        ctx.transposed = transposed
        ctx.dims = (M, N, K)
        ctx.save_for_backward(x2, w, saved_act)
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w, saved_act = ctx.saved_tensors
        M, N, K = ctx.dims
        tr = ctx.transposed
        pre_colsum = take_colsum(dy, N) if ctx.activation is None else None
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        if dy2.dtype != x2.dtype:
            dy2 = dy2.to(x2.dtype)
        if ctx.activation == "gelu":
            dy2 = act_op(dy2, saved_act, 1)
        elif ctx.activation == "relu":
            dy2 = act_op(dy2, saved_act, 2)
        dx = dw = db = None
        weight = ctx.weight_ref
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dY[M,N] . W ; contraction over N: W[in,out] is K-major for it, W[out,in] is MN-major
            dx = gemm(dy2, w, a_mn=False, b_mn=tr, M=M, N=K, K=N).reshape(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            sink = getattr(weight, "_dfb_grad_sink", None)      # engine: write straight into the flat gradient bucket
            if tr:      # dW[out,in] = dY^T . X
                args = dict(a=dy2, b=x2, M=N, N=K)
            else:       # dW[in,out] = X^T . dY
                args = dict(a=x2, b=dy2, M=K, N=N)
            if sink is not None:
                gemm(args["a"], args["b"], a_mn=True, b_mn=True, M=args["M"], N=args["N"], K=M,
                     out=sink.view(args["M"], args["N"]))
                if getattr(weight, "_dfb_grad_parts", 1) > 1:   # tied weight: another producer adds on top of this (ops/embed.py)
                    weight._dfb_sink_filled = True
                weight._dfb_grad_ready()
            else:
                dw = gemm(args["a"], args["b"], a_mn=True, b_mn=True, M=args["M"], N=args["N"], K=M, out_dtype=weight.dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            bias = ctx.bias_ref
            db = pre_colsum if pre_colsum is not None else colsum(dy2)
            sink = getattr(bias, "_dfb_grad_sink", None)
            if sink is not None:
                sink.copy_(db)
                bias._dfb_grad_ready()
                db = None
            else:
                db = db.to(bias.dtype)
        return dx, dw, db, None, None


def linear_cuda(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], activation: Optional[str],
                transposed: bool = False) -> torch.Tensor:
    if activation not in ACT:
        raise ValueError(f"unsupported fused activation {activation}")
    return _LinearFn.apply(x, weight, bias, activation, transposed)


def _deliver(param, grad: torch.Tensor, shape=None):
    """Write a parameter gradient into the engine's bucket sink if there is one (returns None then), else return it."""
    sink = getattr(param, "_dfb_grad_sink", None)
    if sink is not None:
        sink.copy_(grad.reshape(-1))
        param._dfb_grad_ready()
        return None
    return grad.to(param.dtype)


class _MlpFn(torch.autograd.Function):
    """y = gelu(x @ w1 + b1) @ w2 + b2 with the GELU living in GEMM epilogues: forward fc1 writes the pre-activation and the
    GELU output in one pass; backward, fc2's data-gradient GEMM multiplies by gelu'(pre) in its epilogue."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2):
        _init_sigs()
        K, H = w1.shape
        N = w2.shape[1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        pre = torch.empty((M, H), device=x.device, dtype=torch.bfloat16)
        h = gemm_aux(x2, w1, a_mn=False, b_mn=True, M=M, N=H, K=K, aux=pre, aux_mode=1, bias=b1, act="gelu")
        y = gemm(h, w2, a_mn=False, b_mn=True, M=M, N=N, K=H, bias=b2)
        ctx.save_for_backward(x2, pre, h, w1, w2)
        ctx.refs = (w1, b1, w2, b2)
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, pre, h, w1d, w2d = ctx.saved_tensors
        w1, b1, w2, b2 = ctx.refs
        M, K = x2.shape
        H, N = w2d.shape
        pre_colsum = take_colsum(dy, N)
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        # Order matters under the overlapped engine: a parameter's `_dfb_grad_ready()` may launch its bucket's fused
        # reduce-scatter + optimizer + all-gather kernel on the side stream, which OVERWRITES the working weights.  Every
        # kernel that still reads a weight (the data-gradient GEMMs) is therefore enqueued before that weight is reported.
        # fc2: dpre = (dy W2^T) * gelu'(pre)   <- GELU backward in the dgrad epilogue;  dW2 = h^T dy, db2 = colsum(dy)
        # ... and db1 = colsum(dpre) out of that same epilogue
        dpre, dpre_colsum = gemm_dgelu_colsum(dy2, w2d, a_mn=False, b_mn=False, M=M, N=H, K=N, aux=pre) if b1 is not None else \
            (gemm_aux(dy2, w2d, a_mn=False, b_mn=False, M=M, N=H, K=N, aux=pre, aux_mode=2), None)
        dx = gemm(dpre, w1d, a_mn=False, b_mn=False, M=M, N=K, K=H).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw2 = _wgrad(h, dy2, w2, H, N, M)
        db2 = _deliver(b2, pre_colsum if pre_colsum is not None else colsum(dy2)) if b2 is not None else None
        # fc1
        dw1 = _wgrad(x2, dpre, w1, K, H, M)
        db1 = _deliver(b1, dpre_colsum if dpre_colsum is not None else colsum(dpre)) if b1 is not None else None
        return dx, dw1, db1, dw2, db2


def _wgrad(x2, dy2, weight, K, N, M):
    sink = getattr(weight, "_dfb_grad_sink", None)
    if sink is not None:
        gemm(x2, dy2, a_mn=True, b_mn=True, M=K, N=N, K=M, out=sink.view(K, N))
        weight._dfb_grad_ready()
        return None
    return gemm(x2, dy2, a_mn=True, b_mn=True, M=K, N=N, K=M, out_dtype=weight.dtype)


import os as _os

# GELU in the GEMM epilogues (forward: pre-activation and activation stored in one pass; backward: gelu'(pre) multiplied into
# the data-gradient).  `DFB_FUSE_GELU=0/1` overrides for A/B runs (scripts/bench_aux.py).
FUSE_GELU = _os.environ.get("DFB_FUSE_GELU", "1") == "1"
FUSE_DGELU_COLSUM = _os.environ.get("DFB_FUSE_DGELU_COLSUM", "1") == "1"     # fc1 bias gradient out of the gelu'-epilogue GEMM


def mlp_cuda(x, w1, b1, w2, b2):
    ok = FUSE_GELU and x.dtype == torch.bfloat16 and all(t.dtype == torch.bfloat16 for t in (w1, w2)) and w1.shape[1] % 8 == 0 \
        and w1.shape[0] % 8 == 0 and w2.shape[1] % 8 == 0 and w1.is_contiguous() and w2.is_contiguous()
    if not ok:
        return linear_cuda(linear_cuda(x, w1, b1, "gelu"), w2, b2, None)
    return _MlpFn.apply(x, w1, b1, w2, b2)
