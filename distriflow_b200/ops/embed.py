"""Embedding lookup whose backward scatters into the engine's gradient bucket (`csrc/elementwise.cu:
embedding_bwd_scatter_kernel`).  Not in the reference (its models start with Flatten / Conv2D,
/root/reference/experiment/mnist/model.json); GPT-2 of BASELINE.json config 5 needs it, and its table is tied to the LM head:
the head's weight-gradient GEMM writes the bucket slice first (`ops/dense.py: _LinearFn.backward` marks it filled), the
token rows are added on top here, and only then is the parameter reported ready (`_dfb_grad_parts = 2`)."""
from __future__ import annotations

import ctypes
import os

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import c_int, c_void_p, check, dt, ptr, stream


# The in-place scatter adds repeated tokens in an arbitrary order (bf16 rounding per addition), so two runs differ in the last
# bits of the table's gradient.  DETERMINISTIC (env DFB_EMBEDDING_DETERMINISTIC=1, or engine key `deterministic: True`)
# takes PyTorch's sorted segment reduction instead and adds its dense result onto the slice: slower (0.4-0.7 ms of the GPT-2
# step), reproducible.  Attention's fp32 dQ reduction order stays order-dependent for sequences above 256 tokens either way.
DETERMINISTIC_ENV = os.environ.get("DFB_EMBEDDING_DETERMINISTIC", "0") == "1"
DETERMINISTIC = DETERMINISTIC_ENV


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_emb_sigs", False):
        L.dfb_embedding_bwd_scatter.restype = c_int
        L.dfb_embedding_bwd_scatter.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_long, c_int, ctypes.c_long, c_int, c_void_p]
        L._dfb_emb_sigs = True
    return L


def scatter_add_rows(dy2: torch.Tensor, idx: torch.Tensor, table_grad: torch.Tensor) -> None:
    """`table_grad[idx[t]] += dy2[t]` in place.  dy2 `[n, D]` contiguous, idx int64 `[n]`, table_grad `[V, D]` contiguous."""
    L = _sigs()
    n, D = dy2.shape
    check(L.dfb_embedding_bwd_scatter(ptr(dy2), ptr(idx), ptr(table_grad), ctypes.c_long(n), D, ctypes.c_long(table_grad.shape[0]),
                                      dt(table_grad), stream()), "dfb_embedding_bwd_scatter")


class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, idx, table):
        ctx.save_for_backward(idx)
        ctx.table_ref = table
        return F.embedding(idx, table)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        table = ctx.table_ref
        V, D = table.shape
        sink = getattr(table, "_dfb_grad_sink", None)
        dy2 = dy.reshape(-1, D)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        flat_idx = idx.reshape(-1)
        if flat_idx.dtype != torch.int64:
            flat_idx = flat_idx.long()
        ok = sink is not None and dy2.dtype == sink.dtype and sink.dtype in (torch.bfloat16, torch.float32) and \
            D % (8 if sink.dtype == torch.bfloat16 else 4) == 0 and dy2.data_ptr() % 16 == 0 and sink.data_ptr() % 16 == 0
        if not ok or DETERMINISTIC:                   # shapes / dtypes the kernel does not take, or a reproducible sum wanted
            g = torch.ops.aten.embedding_dense_backward(dy2, flat_idx, V, -1, False)      # sort + segment reduce: deterministic
            if sink is not None and getattr(table, "_dfb_sink_filled", False):   # on top of the head's gradient, not instead
                table._dfb_sink_filled = False
                sink.view(V, D).add_(g.to(sink.dtype))
                table._dfb_grad_ready()
                return None, None
            return None, g.to(table.dtype)            # accumulated by autograd, delivered by the engine's hook
        grad2 = sink.view(V, D)
        if getattr(table, "_dfb_sink_filled", False):  # the tied LM head wrote its weight gradient into the slice already
            table._dfb_sink_filled = False
        else:
            grad2.zero_()
        scatter_add_rows(dy2, flat_idx, grad2)
        table._dfb_grad_ready()
        return None, None


def embedding_cuda(idx: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    return _EmbeddingFn.apply(idx, table)
