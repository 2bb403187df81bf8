"""Loss and metric functions selectable by name.

Parity: the reference's `lossesMap` (`/root/reference/src/common/utils.ts:19-30`)
maps names to `tf.losses.*`, all with signature `(labels, predictions)` and the
default `SUM_BY_NONZERO_WEIGHTS` reduction (= mean over elements for unit weights).
Unlike the reference, where `fit` ignores the configured loss and always calls
`tf.losses.softmaxCrossEntropy(prediction, y)` with swapped arguments
(`/root/reference/src/common/models.ts:139`), the configured name
here really selects the training loss.  The swapped-argument behaviour is kept
as the explicit mode `"referenceCompat"` for the MNIST parity check.

On CUDA tensors `softmaxCrossEntropy` dispatches to the fused sm_100a kernel
(`ops/softmax_ce.py`); everything here is the plain PyTorch restatement that
also serves as its numerics oracle.
"""
from __future__ import annotations

from typing import Callable, Dict, List

import torch
import torch.nn.functional as F

__all__ = ["LOSSES", "METRICS", "get_loss", "get_metric", "reference_compat_loss",
           "softmax_cross_entropy"]

_EPS = 1e-7


def _mean(t: torch.Tensor) -> torch.Tensor:
    return t.float().mean()


def absolute_difference(labels, preds):
    return _mean((labels.float() - preds.float()).abs())


def compute_weighted_loss(losses, weights=None):
    # tf.losses.computeWeightedLoss(losses, weights): mean of the (weighted) losses
    if weights is None or not isinstance(weights, torch.Tensor) or weights.shape != losses.shape:
        return _mean(losses)
    w = weights.float()
    return (losses.float() * w).sum() / (w != 0).sum().clamp(min=1)


def hinge_loss(labels, preds):
    lab = 2.0 * labels.float() - 1.0
    return _mean(F.relu(1.0 - lab * preds.float()))


def huber_loss(labels, preds, delta: float = 1.0):
    err = (preds.float() - labels.float()).abs()
    quad = torch.clamp(err, max=delta)
    lin = err - quad
    return _mean(0.5 * quad * quad + delta * lin)


def log_loss(labels, preds, eps: float = _EPS):
    l, p = labels.float(), preds.float()
    return _mean(-l * torch.log(p + eps) - (1 - l) * torch.log(1 - p + eps))


def mean_squared_error(labels, preds):
    return _mean((labels.float() - preds.float()) ** 2)


def sigmoid_cross_entropy(labels, logits):
    return _mean(F.binary_cross_entropy_with_logits(logits.float(), labels.float(), reduction="none"))


def softmax_cross_entropy(labels, logits):
    """Standard fused softmax-CE on logits; `labels` one-hot/soft `[B,C]` or class indices `[B]`."""
    if logits.is_cuda:
        from .ops.softmax_ce import softmax_cross_entropy as fused
        return fused(logits, labels)
    if labels.dtype in (torch.int64, torch.int32) and labels.dim() == logits.dim() - 1:
        return F.cross_entropy(logits.float(), labels.long())
    return _mean(-(labels.float() * F.log_softmax(logits.float(), dim=-1)).sum(-1))


def categorical_crossentropy(labels, probs):
    """Keras categoricalCrossentropy on probabilities (model already ends in softmax)."""
    p = probs.float().clamp(_EPS, 1.0 - _EPS)
    return _mean(-(labels.float() * torch.log(p)).sum(-1))


class _RefCompatCE(torch.autograd.Function):
    """tfjs `softmaxCrossEntropy(onehotLabels=prediction, logits=y)` — the reference's swapped call.

    value  = mean_b( -sum_c p_bc * log_softmax(y_b)_c )
    grad_p = dy/B * (p - softmax(y))          ("custom" — what tfjs-core 1.2.x's customGrad returns
                                                for its first argument; stated from memory)
           | dy/B * (-log_softmax(y))         ("true" derivative of the value)
    No gradient flows to y.
    """

    @staticmethod
    def forward(ctx, pred, y, variant):
        logp = F.log_softmax(y.float(), dim=-1)
        ctx.save_for_backward(pred, logp)
        ctx.variant = variant
        return (-(pred.float() * logp).sum(-1)).mean()

    @staticmethod
    def backward(ctx, dy):
        pred, logp = ctx.saved_tensors
        b = pred.shape[0]
        if ctx.variant == "custom":
            g = (pred.float() - logp.exp()) * (dy / b)
        else:
            g = -logp * (dy / b)
        return g.to(pred.dtype), None, None


def reference_compat_loss(labels, preds, variant: str = "custom"):
    """`labels` = dataset one-hot target, `preds` = model output (signature kept `(labels, preds)`)."""
    if preds.is_cuda:
        from .ops.softmax_ce import reference_compat_ce
        return reference_compat_ce(labels, preds, variant)
    return _RefCompatCE.apply(preds, labels, variant)


def reference_compat_true(labels, preds):
    return reference_compat_loss(labels, preds, "true")


LOSSES: Dict[str, Callable] = {
    "absoluteDifference": absolute_difference,
    "computeWeightedLoss": compute_weighted_loss,
    "hingeLoss": hinge_loss,
    "huberLoss": huber_loss,
    "logLoss": log_loss,
    "meanSquaredError": mean_squared_error,
    "sigmoidCrossEntropy": sigmoid_cross_entropy,
    "softmaxCrossEntropy": softmax_cross_entropy,
    # additions
    "categoricalCrossentropy": categorical_crossentropy,
    "referenceCompat": reference_compat_loss,
    "referenceCompatTrueGrad": reference_compat_true,
}


def accuracy(labels, preds) -> torch.Tensor:
    """Keras 'accuracy' for categorical targets: argmax match rate."""
    if labels.dim() == preds.dim() and labels.shape[-1] == preds.shape[-1] and preds.shape[-1] > 1:
        lab = labels.argmax(-1)
    else:
        lab = labels.reshape(preds.shape[:-1]).long()
    return (preds.argmax(-1) == lab).float().mean()


METRICS: Dict[str, Callable] = {"accuracy": accuracy, "acc": accuracy}


def get_loss(name: str) -> Callable:
    if name not in LOSSES:
        raise KeyError(f'unknown loss "{name}"; known: {sorted(LOSSES)}')
    return LOSSES[name]


def get_metric(name: str) -> Callable:
    if name in METRICS:
        return METRICS[name]
    if name in LOSSES:
        return LOSSES[name]
    raise KeyError(f'unknown metric "{name}"')


def metric_names(metrics: List[str]) -> List[str]:
    return list(metrics)
