"""PyTorch DDP + NCCL (+ cuBLAS/cuDNN) baseline of the identical model and schedule (BASELINE.json "Baselines").

The model is the same module as in the product arm but every hot op is forced onto the stock PyTorch/cuBLAS
kernels (`F.linear`, `F.conv2d`, `F.cross_entropy`), gradients are averaged by `DistributedDataParallel`'s
bucketed NCCL all-reduce overlapped with backward, and the optimizer is `torch.optim` (fused).  Weights and
gradients are bf16 like the product arm's working copies, with the same bucket size.
"""
from __future__ import annotations

import contextlib

_STOCK_CTX = None

import torch
import torch.nn.functional as F


@contextlib.contextmanager
def stock_ops():
    """Route distriflow_b200.ops through PyTorch library kernels (baseline arm only)."""
    from distriflow_b200 import ops
    saved = (ops.linear, ops.conv2d_nhwc, ops.maxpool2d_nhwc, ops.softmax_cross_entropy, ops.batch_norm_nhwc)
    from distriflow_b200.ops.norm import batch_norm_reference

    def bn(x, gamma, beta, rm, rv, training, momentum=0.1, eps=1e-5, activation=None, residual=None):
        return batch_norm_reference(x, gamma, beta, rm, rv, training, momentum, eps, activation, residual)

    def linear(x, w, b=None, activation=None, transposed=False):
        y = F.linear(x, w if transposed else w.t(), b)
        if activation == "relu":
            y = F.relu(y)
        elif activation == "gelu":
            y = F.gelu(y, approximate="tanh")
        return y

    def conv(x, k, b=None, strides=(1, 1), padding="valid", activation=None):
        xc = x.permute(0, 3, 1, 2)
        kh, kw = k.shape[0], k.shape[1]
        if padding == "same":
            pt, pb = ops.same_padding(x.shape[1], kh, strides[0])
            pl, pr = ops.same_padding(x.shape[2], kw, strides[1])
            xc = F.pad(xc, (pl, pr, pt, pb))
        y = F.conv2d(xc, k.permute(3, 2, 0, 1), b, stride=tuple(strides)).permute(0, 2, 3, 1)
        return F.relu(y) if activation == "relu" else y

    def pool(x, pool=(2, 2), strides=None, padding="valid"):
        return F.max_pool2d(x.permute(0, 3, 1, 2), tuple(pool), tuple(strides or pool)).permute(0, 2, 3, 1)

    def ce(logits, labels, ignore_index=None, inplace=False, unit_grad=False):
        if labels.dtype in (torch.int64, torch.int32) and labels.dim() == logits.dim() - 1:
            return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), labels.reshape(-1).long())
        return (-(labels.float() * F.log_softmax(logits.float(), -1)).sum(-1)).mean()

    ops.linear, ops.conv2d_nhwc, ops.maxpool2d_nhwc, ops.softmax_cross_entropy, ops.batch_norm_nhwc = linear, conv, pool, ce, bn
    try:
        yield
    finally:
        ops.linear, ops.conv2d_nhwc, ops.maxpool2d_nhwc, ops.softmax_cross_entropy, ops.batch_norm_nhwc = saved


def make_ddp_step(model, loss_fn, optimizer: str, lr: float, world: int, device, bucket_mb: int = 32):
    model = model.to(device).to(torch.bfloat16)
    ddp = model
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP
        ddp = DDP(model, device_ids=[device.index], bucket_cap_mb=bucket_mb, gradient_as_bucket_view=True)
    params = [p for p in model.parameters() if p.requires_grad]
    if optimizer == "adam":
        opt = torch.optim.Adam(params, lr=lr, fused=True)
    elif optimizer == "momentum":
        opt = torch.optim.SGD(params, lr=lr, momentum=0.9)
    else:
        opt = torch.optim.SGD(params, lr=lr)
    global _STOCK_CTX
    _STOCK_CTX = stock_ops()        # kept referenced: a collected generator would run its `finally` and undo the patch
    _STOCK_CTX.__enter__()          # stays on for the life of the process (baseline arm)

    def step(*batch):
        opt.zero_grad(set_to_none=True)
        if world > 1:
            loss = loss_fn(ddp.module if False else _Wrap(ddp, model), *batch)
        else:
            loss = loss_fn(model, *batch)
        loss.backward()
        opt.step()
        return loss.detach()

    return step


class _Wrap:
    """Lets `loss_fn(model, *batch)` call through the DDP wrapper (so its hooks fire) while exposing attributes."""

    def __init__(self, ddp, inner):
        self._ddp, self._inner = ddp, inner

    def __call__(self, *a, **k):
        return self._ddp(*a, **k)

    def __getattr__(self, name):
        return getattr(self._inner, name)
