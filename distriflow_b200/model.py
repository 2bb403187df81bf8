"""DistriModel: the model contract shared by servers and workers.

Parity:
  * `DistributedModel` interface — `fit(x,y) -> grads[]` (does NOT change the
    weights), `update(grads)`, `predict`, `evaluate -> number[]`, `getVars`,
    `setVars`, `inputShape`, `outputShape`
    (`/root/reference/src/common/models.ts:7-72`);
  * `DistributedTfModel` — wrapper around a layers model with SGD
    (`models.ts:74-151`) -> `DistriTorchModel` around a `torch.nn.Module`;
  * `DistributedDynamicModel` — functional model from
    `{vars, predict, loss, inputShape, outputShape}` (`models.ts:153-208`)
    -> `DistriDynamicModel` (with the `-lr*g` NaN bug of `models.ts:185` fixed);
  * README design fields `modelID`/`version`, `maximumStaleness`,
    `savedGradient` (`/root/reference/README.md:23-30`), which the reference
    never implemented.

Gradients are ordered by the model's own parameter list (fixes SURVEY Q14).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence

import torch

from .config import DEFAULT_CLIENT_HYPERPARAMS, compile_args
from .losses import get_loss, get_metric

__all__ = ["DistriModel", "DistriTorchModel", "DistriDynamicModel", "StalenessPolicy",
           "DistributedModel", "DistributedTfModel", "DistributedDynamicModel"]


class StalenessPolicy:
    """Bounded-staleness rule: `staleness = serverVersionIndex - gradientVersionIndex`.

    `maximumStaleness=0` reproduces the federated server's "current version only" rule
    (`/root/reference/src/server/federated_server.ts:73`); the async server of the
    reference applies everything (`asynchronousSGD_server.ts:65-73`) — use a large bound
    for that behaviour."""

    def __init__(self, maximum_staleness: int = 0, policy: str = "drop"):
        self.maximumStaleness = int(maximum_staleness)
        self.policy = policy

    def weight(self, staleness: int) -> float:
        """0.0 = reject; otherwise the factor applied to the gradient."""
        if staleness < 0 or staleness > self.maximumStaleness:
            return 0.0
        if self.policy == "scale":
            return 1.0 / (1.0 + staleness)
        return 1.0


class DistriModel:
    """Abstract contract + README bookkeeping (version, savedGradient, maximumStaleness)."""

    isDistributedModel = True

    def __init__(self) -> None:
        self.version: str = "0"
        self.versionIndex: int = 0          # monotone counter used for staleness arithmetic
        self.maximumStaleness: int = 0
        self.savedGradient: Optional[List[torch.Tensor]] = None
        self.savedCount: int = 0

    # -- contract -----------------------------------------------------------------------------
    def fit(self, x, y) -> List[torch.Tensor]:
        raise NotImplementedError

    def update(self, grads: Sequence[torch.Tensor]) -> None:
        raise NotImplementedError

    def predict(self, x) -> torch.Tensor:
        raise NotImplementedError

    def evaluate(self, x, y) -> List[float]:
        raise NotImplementedError

    def get_vars(self) -> List[torch.Tensor]:
        raise NotImplementedError

    def set_vars(self, vals: Sequence[torch.Tensor]) -> None:
        raise NotImplementedError

    inputShape: Sequence[int] = ()
    outputShape: Sequence[int] = ()

    def setup(self) -> None:
        """Resolve / compile the model (reference: `fetchInitial`)."""

    # reference spellings
    def getVars(self):
        return self.get_vars()

    def setVars(self, vals):
        return self.set_vars(vals)

    @property
    def modelID(self) -> str:
        return self.version

    # -- savedGradient ------------------------------------------------------------------------
    def accumulate(self, grads: Sequence[torch.Tensor], weight: float = 1.0) -> int:
        """Add one worker's gradient into `savedGradient`; returns the number buffered."""
        if self.savedGradient is None:
            self.savedGradient = [g.detach().float().clone() * weight for g in grads]
        else:
            if len(grads) != len(self.savedGradient):
                raise ValueError("gradient list length mismatch")
            for acc, g in zip(self.savedGradient, grads):
                acc.add_(g.detach().to(acc.device, torch.float32), alpha=weight)
        self.savedCount += 1
        return self.savedCount

    def apply_saved(self, aggregation: str = "mean") -> int:
        """Reduce `savedGradient` (mean over contributions) and apply it with `update`."""
        if self.savedGradient is None:
            return 0
        if aggregation != "mean":           # federated_server.ts:104
            raise ValueError(f"unsupported aggregation {aggregation}")
        n = self.savedCount
        grads = [g / n for g in self.savedGradient]
        self.update(grads)
        self.savedGradient, self.savedCount = None, 0
        return n


class DistriTorchModel(DistriModel):
    """`torch.nn.Module` wrapper; optimiser is plain SGD as in the reference (`models.ts:87-89,128-135`)."""

    def __init__(self, initial_model: Any, config: Optional[Dict[str, Any]] = None,
                 device: Optional[torch.device] = None):
        super().__init__()
        cfg = compile_args(config)
        self._initial = initial_model
        self.loss = cfg["loss"]
        self.lossFunction = get_loss(self.loss)
        self.metrics = list(cfg["metrics"])
        self.optimizer = "sgd"
        self.learningRate = float(cfg["learningRate"])
        self.device = device
        self.model: Optional[torch.nn.Module] = None
        if isinstance(initial_model, torch.nn.Module):
            self._resolve()

    # -- model fetch (utils.ts:236-244) -------------------------------------------------------
    def _resolve(self) -> None:
        from .checkpoint import fetch_model
        if self.model is None:
            self.model = fetch_model(self._initial)
            if self.device is not None:
                self.model.to(self.device)
            self._params = [p for p in self.model.parameters() if p.requires_grad]

    def fetch_initial(self) -> None:
        self._resolve()

    fetchInitial = fetch_initial

    def setup(self) -> None:
        self._resolve()

    # -- contract -----------------------------------------------------------------------------
    def _dev(self, t):
        if isinstance(t, torch.Tensor) and self._params and t.device != self._params[0].device:
            return t.to(self._params[0].device, non_blocking=True)
        return t

    def predict(self, x):
        self._resolve()
        with torch.no_grad():
            return self.model(self._dev(x))

    def fit(self, x, y) -> List[torch.Tensor]:
        """Forward + backward on one micro-batch; returns gradients, leaves weights untouched."""
        self._resolve()
        x, y = self._dev(x), self._dev(y)
        self.model.train()
        out = self.model(x)
        loss = self.lossFunction(y, out)
        self.lastLoss = loss.detach()
        grads = torch.autograd.grad(loss, self._params, allow_unused=True)
        return [g if g is not None else torch.zeros_like(p) for g, p in zip(grads, self._params)]

    def update(self, grads: Sequence[torch.Tensor]) -> None:
        self._resolve()
        with torch.no_grad():
            for p, g in zip(self._params, grads):
                p.add_(g.to(p.device, p.dtype), alpha=-self.learningRate)

    def evaluate(self, x, y) -> List[float]:
        """[compiled loss, *metrics] on the batch (`models.ts:106-115`)."""
        self._resolve()
        x, y = self._dev(x), self._dev(y)
        was_training = self.model.training
        self.model.eval()
        with torch.no_grad():
            out = self.model(x)
            vals = [self.lossFunction(y, out)] + [get_metric(m)(y, out) for m in self.metrics]
        self.model.train(was_training)
        return [float(v) for v in vals]

    def get_vars(self) -> List[torch.Tensor]:
        self._resolve()
        return [p.data for p in self._params]

    def set_vars(self, vals: Sequence[torch.Tensor]) -> None:
        self._resolve()
        if len(vals) != len(self._params):
            raise ValueError(f"expected {len(self._params)} variables, got {len(vals)}")
        with torch.no_grad():
            for p, v in zip(self._params, vals):
                if tuple(p.shape) != tuple(v.shape):
                    raise ValueError(f"shape mismatch: {tuple(p.shape)} vs {tuple(v.shape)}")
                p.copy_(v.to(p.device, p.dtype))

    @property
    def inputShape(self):
        self._resolve()
        return list(getattr(self.model, "input_shape", ()))

    @property
    def outputShape(self):
        self._resolve()
        return list(getattr(self.model, "output_shape", ()))


class DistriDynamicModel(DistriModel):
    """Functional model: user supplies variables, `predict(x)` and `loss(labels, preds)`."""

    isDistributedClientModel = True

    def __init__(self, *, vars: Sequence[torch.Tensor], predict: Callable, loss: Callable,
                 inputShape: Sequence[int], outputShape: Sequence[int],
                 learningRate: float = DEFAULT_CLIENT_HYPERPARAMS["learningRate"]):
        super().__init__()
        self.vars = [v if v.requires_grad else v.requires_grad_(True) for v in vars]
        self._predict = predict
        self.loss = loss
        self.learningRate = float(learningRate)
        self._in, self._out = list(inputShape), list(outputShape)

    @property
    def inputShape(self):
        return self._in

    @property
    def outputShape(self):
        return self._out

    def predict(self, x):
        with torch.no_grad():
            return self._predict(x)

    def fit(self, x, y):
        loss = self.loss(y, self._predict(x)).mean()
        self.lastLoss = loss.detach()
        grads = torch.autograd.grad(loss, self.vars, allow_unused=True)
        return [g if g is not None else torch.zeros_like(v) for g, v in zip(grads, self.vars)]

    def update(self, grads):
        with torch.no_grad():
            for v, g in zip(self.vars, grads):
                v.add_(g.to(v.device, v.dtype), alpha=-self.learningRate)   # v -= lr*g (Q12 fixed)

    def evaluate(self, x, y):
        with torch.no_grad():
            return [float(v) for v in self.loss(y, self._predict(x)).reshape(-1)]

    def get_vars(self):
        return [v.data for v in self.vars]

    def set_vars(self, vals):
        with torch.no_grad():
            for v, val in zip(self.vars, vals):
                v.copy_(val.to(v.device, v.dtype))


# reference class names
DistributedModel = DistriModel
DistributedTfModel = DistriTorchModel
DistributedDynamicModel = DistriDynamicModel
