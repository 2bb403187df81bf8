"""PyTorch DDP + NCCL (+ cuBLAS/cuDNN) baseline of the identical model and schedule (BASELINE.json "Baselines").

The model is the same module as in the product arm but every op of `distriflow_b200.ops` is forced onto the stock
PyTorch/cuBLAS/cuDNN kernels (`F.linear`, `F.conv2d`, `F.cross_entropy`, `F.layer_norm`, `F.gelu`, SDPA; none of this
repo's kernels run in this arm), gradients are averaged by `DistributedDataParallel`'s
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
    names = ("linear", "conv2d_nhwc", "maxpool2d_nhwc", "softmax_cross_entropy", "batch_norm_nhwc", "mlp", "layer_norm",
             "add_layer_norm", "causal_self_attention", "embedding")
    saved = tuple(getattr(ops, n) for n in names)
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

    def mlp(x, w1, b1, w2, b2):
        return F.linear(F.gelu(F.linear(x, w1.t(), b1), approximate="tanh"), w2.t(), b2)

    def layer_norm(x, gamma, beta, eps=1e-5):
        return F.layer_norm(x, (x.shape[-1],), gamma.to(x.dtype), beta.to(x.dtype), eps)

    def add_layer_norm(x, r, gamma, beta, eps=1e-5):
        s = x + r
        return s, F.layer_norm(s, (s.shape[-1],), gamma.to(s.dtype), beta.to(s.dtype), eps)

    def attention(qkv, n_head):
        B, T, C3 = qkv.shape
        D = C3 // (3 * n_head)
        q, k, v = qkv.view(B, T, 3, n_head, D).permute(2, 0, 3, 1, 4)
        return F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, n_head * D)

    def embedding(idx, table):
        return F.embedding(idx, table)

    for n, f in zip(names, (linear, conv, pool, ce, bn, mlp, layer_norm, add_layer_norm, attention, embedding)):
        setattr(ops, n, f)
    try:
        yield
    finally:
        for n, f in zip(names, saved):
            setattr(ops, n, f)


def make_ddp_step(model, loss_fn, optimizer: str, lr: float, world: int, device, bucket_mb: int = 32, compile_model: bool = False):
    """Returns `(step, close)`; `close()` restores `distriflow_b200.ops` (the patch is process-wide while the arm runs)."""
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
    _STOCK_CTX = ctx = stock_ops()  # kept referenced: a collected generator would run its `finally` and undo the patch
    ctx.__enter__()
    target = _Wrap(ddp, model) if world > 1 else model

    def fwd_loss(*batch):
        return loss_fn(target, *batch)

    run = torch.compile(fwd_loss) if compile_model else fwd_loss     # stronger variant of the baseline (never of the product)

    def step(*batch):
        opt.zero_grad(set_to_none=True)
        loss = run(*batch)
        loss.backward()
        opt.step()
        return loss.detach()

    def close():
        global _STOCK_CTX
        ctx.__exit__(None, None, None)
        _STOCK_CTX = None

    return step, close


class _Wrap:
    """Lets `loss_fn(model, *batch)` call through the DDP wrapper (so its hooks fire) while exposing attributes."""

    def __init__(self, ddp, inner):
        self._ddp, self._inner = ddp, inner

    def __call__(self, *a, **k):
        return self._ddp(*a, **k)

    def __getattr__(self, name):
        return getattr(self._inner, name)
