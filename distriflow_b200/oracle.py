"""CPU oracle: a from-scratch restatement of the reference's tf.js math and training schedules.

The reference cannot run here (no node, no tf.js), so parity is claimed against this module:
NumPy fp32 forward/backward of the MNIST dense model
(`/root/reference/experiment/mnist/mnist_server.ts:16-22`), the swapped-argument loss of
`/root/reference/src/common/models.ts:139` with both gradient variants, plain SGD
(`models.ts:128-135`), and single-process simulators of the two schedules:
  * `simulate_async`  — `/root/reference/src/server/asynchronousSGD_server.ts:65-80` with the intended
    FCFS-per-worker dispatch (Q6) and an optional staleness bound (Q7);
  * `simulate_sync`   — `/root/reference/src/server/federated_server.ts:71-116` with
    `minUpdatesPerVersion = #workers` (synchronous batch gradient descent).
It is deliberately independent of `torch.autograd` and of this package's ops so it can serve as their check,
and it is the timed CPU baseline of BASELINE.md.
"""
from __future__ import annotations

import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

__all__ = ["MlpOracle", "CnnOracle", "log_softmax", "softmax", "ref_compat_loss_and_grad", "softmax_ce_loss_and_grad",
           "simulate_async", "simulate_sync", "synthetic_mnist"]


def log_softmax(z: np.ndarray) -> np.ndarray:
    m = z.max(-1, keepdims=True)
    return z - m - np.log(np.exp(z - m).sum(-1, keepdims=True))


def softmax(z: np.ndarray) -> np.ndarray:
    return np.exp(log_softmax(z))


def ref_compat_loss_and_grad(pred: np.ndarray, y: np.ndarray, variant: str = "custom") -> Tuple[float, np.ndarray]:
    """tfjs `softmaxCrossEntropy(onehotLabels=pred, logits=y).mean()` and d/dpred.

    custom: `dy*(labels - softmax(logits))` — the customGrad of tfjs-core 1.2.x for its first argument.
    true  : `-log_softmax(y)` — the derivative of the value itself."""
    b = pred.shape[0]
    lp = log_softmax(y.astype(np.float32))
    loss = float((-(pred * lp).sum(-1)).mean())
    g = (pred - np.exp(lp)) / b if variant == "custom" else -lp / b
    return loss, g.astype(np.float32)


def softmax_ce_loss_and_grad(logits: np.ndarray, y: np.ndarray) -> Tuple[float, np.ndarray]:
    b = logits.shape[0]
    lp = log_softmax(logits.astype(np.float32))
    loss = float((-(y * lp).sum(-1)).mean())
    return loss, ((np.exp(lp) * y.sum(-1, keepdims=True) - y) / b).astype(np.float32)


class MlpOracle:
    """Flatten -> Dense(hidden, relu) -> Dense(classes, softmax); weights `[in,out]` like tf.js."""

    def __init__(self, in_features: int = 784, hidden: int = 10, classes: int = 10, seed: int = 0,
                 loss: str = "referenceCompat", lr: float = 0.001):
        rng = np.random.default_rng(seed)

        def glorot(i, o):
            lim = np.sqrt(6.0 / (i + o))
            return rng.uniform(-lim, lim, size=(i, o)).astype(np.float32)

        self.vars: List[np.ndarray] = [glorot(in_features, hidden), np.zeros(hidden, np.float32),
                                       glorot(hidden, classes), np.zeros(classes, np.float32)]
        self.loss, self.lr = loss, lr

    def get_vars(self) -> List[np.ndarray]:
        return [v.copy() for v in self.vars]

    def set_vars(self, vs: Sequence[np.ndarray]) -> None:
        self.vars = [np.asarray(v, np.float32).copy() for v in vs]

    def forward(self, x: np.ndarray):
        w1, b1, w2, b2 = self.vars
        x2 = x.reshape(x.shape[0], -1).astype(np.float32)
        h_pre = x2 @ w1 + b1
        h = np.maximum(h_pre, 0)
        z = h @ w2 + b2
        return x2, h_pre, h, z, softmax(z)

    def predict(self, x: np.ndarray) -> np.ndarray:
        return self.forward(x)[-1]

    def fit(self, x: np.ndarray, y: np.ndarray) -> Tuple[float, List[np.ndarray]]:
        """Loss and gradients for one micro-batch; weights untouched (`models.ts:137-142`)."""
        w1, b1, w2, b2 = self.vars
        x2, h_pre, h, z, p = self.forward(x)
        y = y.astype(np.float32)
        loss, dz = _loss_grad_wrt_logits(self.loss, z, p, y)
        dw2, db2 = h.T @ dz, dz.sum(0)
        dh = (dz @ w2.T) * (h_pre > 0)
        dw1, db1 = x2.T @ dh, dh.sum(0)
        return loss, [g.astype(np.float32) for g in (dw1, db1, dw2, db2)]

    def update(self, grads: Sequence[np.ndarray], lr: Optional[float] = None) -> None:
        lr = self.lr if lr is None else lr
        self.vars = [v - np.float32(lr) * g for v, g in zip(self.vars, grads)]

    def evaluate(self, x: np.ndarray, y: np.ndarray) -> List[float]:
        p = self.predict(x)
        return [float(((y - p) ** 2).mean()), float((p.argmax(-1) == y.argmax(-1)).mean())]


def _loss_grad_wrt_logits(loss_name: str, z: np.ndarray, p: np.ndarray, y: np.ndarray) -> Tuple[float, np.ndarray]:
    """Loss value and d loss / d logits for a model that ends in a softmax layer (shared by both oracles)."""
    if loss_name in ("referenceCompat", "referenceCompatTrueGrad"):
        loss, dp = ref_compat_loss_and_grad(p, y, "custom" if loss_name == "referenceCompat" else "true")
        return loss, p * (dp - (dp * p).sum(-1, keepdims=True))      # back through the model's own softmax layer
    if loss_name == "softmaxCrossEntropy":                           # standard CE on the logits
        return softmax_ce_loss_and_grad(z, y)
    if loss_name == "categoricalCrossentropy":
        loss = float((-(y * np.log(np.clip(p, 1e-7, 1 - 1e-7))).sum(-1)).mean())
        return loss, (p * y.sum(-1, keepdims=True) - y) / z.shape[0]
    if loss_name == "meanSquaredError":
        dp = 2 * (p - y) / p.size
        return float(((y - p) ** 2).mean()), p * (dp - (dp * p).sum(-1, keepdims=True))
    raise ValueError(loss_name)


def _conv_valid(x: np.ndarray, w: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """NHWC x HWIO valid convolution, stride 1.  Returns (y, patch matrix) — K2 of SURVEY §2.5."""
    kh, kw, c, k = w.shape
    b, h, ww, _ = x.shape
    oh, ow = h - kh + 1, ww - kw + 1
    win = np.lib.stride_tricks.sliding_window_view(x, (kh, kw), axis=(1, 2))          # [B, OH, OW, C, kh, kw]
    patches = np.ascontiguousarray(win.transpose(0, 1, 2, 4, 5, 3)).reshape(b * oh * ow, kh * kw * c)
    return (patches @ w.reshape(kh * kw * c, k)).reshape(b, oh, ow, k), patches


def _conv_valid_dx(dy: np.ndarray, w: np.ndarray) -> np.ndarray:
    """Data gradient of `_conv_valid`: full correlation of dy with the flipped filter, channels swapped."""
    kh, kw, c, k = w.shape
    dyp = np.pad(dy, ((0, 0), (kh - 1, kh - 1), (kw - 1, kw - 1), (0, 0)))
    wf = np.ascontiguousarray(w[::-1, ::-1].transpose(0, 1, 3, 2))                      # [kh, kw, K, C]
    return _conv_valid(dyp, wf)[0]


class CnnOracle:
    """The Keras-converted MNIST CNN of `/root/reference/experiment/mnist/model.json`, NumPy fp32, NHWC / HWIO:
    Conv2D(32,3x3,valid)-ReLU-Conv2D(32,3x3,valid)-ReLU-MaxPool 2x2-Flatten(H,W,C)-Dense(128)-ReLU-Dense(classes)-softmax.
    Dropout is inactive, as in the reference's `predictOnBatch`.  K1-K6 for the CNN config."""

    def __init__(self, classes: int = 5, seed: int = 0, loss: str = "referenceCompat", lr: float = 0.001):
        rng = np.random.default_rng(seed)

        def glorot(shape, fan_in, fan_out):
            lim = np.sqrt(6.0 / (fan_in + fan_out))
            return rng.uniform(-lim, lim, size=shape).astype(np.float32)

        self.vars: List[np.ndarray] = [
            glorot((3, 3, 1, 32), 9, 288), np.zeros(32, np.float32),
            glorot((3, 3, 32, 32), 288, 288), np.zeros(32, np.float32),
            glorot((4608, 128), 4608, 128), np.zeros(128, np.float32),
            glorot((128, classes), 128, classes), np.zeros(classes, np.float32)]
        self.loss, self.lr = loss, lr

    get_vars, set_vars, update = MlpOracle.get_vars, MlpOracle.set_vars, MlpOracle.update

    def forward(self, x: np.ndarray):
        k1, b1, k2, b2, w3, b3, w4, b4 = self.vars
        x = x.astype(np.float32).reshape(x.shape[0], 28, 28, 1)
        a1_pre, p1 = _conv_valid(x, k1)
        a1_pre = a1_pre + b1
        a1 = np.maximum(a1_pre, 0)
        a2_pre, p2 = _conv_valid(a1, k2)
        a2_pre = a2_pre + b2
        a2 = np.maximum(a2_pre, 0)
        bsz = x.shape[0]
        win = a2.reshape(bsz, 12, 2, 12, 2, 32)
        pool = win.max(axis=(2, 4))                                                    # [B, 12, 12, 32]
        flat = pool.reshape(bsz, 4608)                                                 # H, W, C order
        h_pre = flat @ w3 + b3
        h = np.maximum(h_pre, 0)
        z = h @ w4 + b4
        return dict(x=x, p1=p1, a1_pre=a1_pre, a1=a1, p2=p2, a2_pre=a2_pre, a2=a2, pool=pool, flat=flat, h_pre=h_pre, h=h,
                    z=z, p=softmax(z))

    def predict(self, x: np.ndarray) -> np.ndarray:
        return self.forward(x)["p"]

    def fit(self, x: np.ndarray, y: np.ndarray) -> Tuple[float, List[np.ndarray]]:
        k1, b1, k2, b2, w3, b3, w4, b4 = self.vars
        f = self.forward(x)
        loss, dz = _loss_grad_wrt_logits(self.loss, f["z"], f["p"], y.astype(np.float32))
        dw4, db4 = f["h"].T @ dz, dz.sum(0)
        dh = (dz @ w4.T) * (f["h_pre"] > 0)
        dw3, db3 = f["flat"].T @ dh, dh.sum(0)
        bsz = f["x"].shape[0]
        dpool = (dh @ w3.T).reshape(bsz, 12, 12, 32)
        # max-pool backward: the gradient goes to the window maximum (first one on a tie; ties are zeros behind a dead ReLU)
        win = f["a2"].reshape(bsz, 12, 2, 12, 2, 32)
        mask = win == f["pool"][:, :, None, :, None, :]
        first = np.cumsum(np.cumsum(mask, axis=2), axis=4) == 1
        da2 = ((mask & first) * dpool[:, :, None, :, None, :]).reshape(bsz, 24, 24, 32)
        da2 = da2 * (f["a2_pre"] > 0)
        dk2 = (f["p2"].T @ da2.reshape(-1, 32)).reshape(3, 3, 32, 32)
        db2 = da2.sum((0, 1, 2))
        da1 = _conv_valid_dx(da2, k2) * (f["a1_pre"] > 0)
        dk1 = (f["p1"].T @ da1.reshape(-1, 32)).reshape(3, 3, 1, 32)
        db1 = da1.sum((0, 1, 2))
        return loss, [g.astype(np.float32) for g in (dk1, db1, dk2, db2, dw3, db3, dw4, db4)]

    def evaluate(self, x: np.ndarray, y: np.ndarray) -> List[float]:
        p = self.predict(x)
        return [float(((y - p) ** 2).mean()), float((p.argmax(-1) == y.argmax(-1)).mean())]


def synthetic_mnist(n: int = 512, classes: int = 10, seed: int = 0, raw_0_255: bool = True):
    """Synthetic MNIST-shaped data `[n,28,28,1]` float32 in 0..255 (Q4) with one-hot int labels: class-dependent
    blobs, so a model can actually reduce the loss."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, classes, size=n)
    protos = rng.uniform(0, 1, size=(classes, 28, 28, 1)).astype(np.float32)
    x = protos[labels] * 0.7 + rng.uniform(0, 0.3, size=(n, 28, 28, 1)).astype(np.float32)
    if raw_0_255:
        x = np.floor(x * 255.0).astype(np.float32)
    y = np.eye(classes, dtype=np.float32)[labels]
    return x, y


def _batches(n: int, bs: int) -> int:
    return n // bs


def simulate_async(model: MlpOracle, x: np.ndarray, y: np.ndarray, *, batch_size: int = 32, epochs: int = 1,
                   workers: int = 1, maximum_staleness: Optional[int] = None,
                   order: Optional[Callable[[int], int]] = None) -> Dict[str, list]:
    """Round-robin (or `order(step) -> worker`) completion order; every worker holds (weights snapshot, batch).

    Returns per-applied-update `loss` (the worker's training loss on its stale weights), `staleness`, `dropped`."""
    nb = _batches(x.shape[0], batch_size)
    queue = [(e, b) for e in range(epochs) for b in range(nb)]
    qi = 0
    version = 0
    hold: List[Optional[Tuple[List[np.ndarray], int, Tuple[int, int]]]] = [None] * workers
    for w in range(workers):
        if qi < len(queue):
            hold[w] = (model.get_vars(), version, queue[qi])
            qi += 1
    out = {"loss": [], "staleness": [], "dropped": 0, "versions": 0, "initial_vars": model.get_vars()}
    step = 0
    retry: List[Tuple[int, int]] = []
    while any(h is not None for h in hold):
        w = order(step) % workers if order else step % workers
        step += 1
        if hold[w] is None:
            continue
        snap, v, (e, b) = hold[w]
        server_vars = model.get_vars()
        model.set_vars(snap)
        loss, grads = model.fit(x[b * batch_size:(b + 1) * batch_size], y[b * batch_size:(b + 1) * batch_size])
        model.set_vars(server_vars)
        staleness = version - v
        if maximum_staleness is not None and staleness > maximum_staleness:
            out["dropped"] += 1
            retry.append((e, b))
        else:
            model.update(grads)
            version += 1
            out["loss"].append(loss)
            out["staleness"].append(staleness)
        nxt = None
        if qi < len(queue):
            nxt = queue[qi]
            qi += 1
        elif retry:
            nxt = retry.pop(0)
        hold[w] = (model.get_vars(), version, nxt) if nxt is not None else None
    out["versions"] = version
    return out


def simulate_sync(model: MlpOracle, x: np.ndarray, y: np.ndarray, *, batch_size: int = 32, epochs: int = 1,
                  workers: int = 2) -> Dict[str, list]:
    """Synchronous batch GD: `workers` micro-batches per version, mean of their gradients, one SGD step."""
    nb = _batches(x.shape[0], batch_size)
    out = {"loss": [], "versions": 0}
    for _ in range(epochs):
        for b0 in range(0, nb - nb % workers if nb >= workers else nb, workers):
            losses, acc = [], None
            group = list(range(b0, min(b0 + workers, nb)))
            for b in group:
                loss, grads = model.fit(x[b * batch_size:(b + 1) * batch_size], y[b * batch_size:(b + 1) * batch_size])
                losses.append(loss)
                acc = grads if acc is None else [a + g for a, g in zip(acc, grads)]
            model.update([a / len(group) for a in acc])
            out["loss"].append(float(np.mean(losses)))
            out["versions"] += 1
    return out


def timed_cpu_baseline(steps: int = 200, batch_size: int = 32, seed: int = 0, model: str = "mlp") -> Dict[str, float]:
    """Samples/s of the oracle on this CPU (the 'timed CPU baseline' of BASELINE.md): `model="mlp"` the MNIST dense config,
    `model="cnn"` the `model.json` CNN topology (inputs / 255, softmax cross-entropy)."""
    if model == "cnn":
        x, y = synthetic_mnist(batch_size * 8, classes=10, seed=seed)
        x = (x / 255.0).astype(np.float32).reshape(-1, 28, 28, 1)
        m: Any = CnnOracle(classes=10, seed=seed, loss="softmaxCrossEntropy")
    else:
        x, y = synthetic_mnist(batch_size * 8, seed=seed)
        m = MlpOracle(seed=seed)
    t0 = time.perf_counter()
    for s in range(steps):
        b = s % 8
        _, g = m.fit(x[b * batch_size:(b + 1) * batch_size], y[b * batch_size:(b + 1) * batch_size])
        m.update(g)
    dt = time.perf_counter() - t0
    return {"samples_per_sec": steps * batch_size / dt, "ms_per_step": dt / steps * 1e3}
