"""Layer modules that keep parameters in the tf.js layouts.

Dense kernels are `[in, out]`, Conv2D kernels are HWIO `[kh, kw, in, out]`,
activations are NHWC (channels-last) end to end, and Flatten after a conv
orders features H,W,C — so a checkpoint in the reference's
`model.json + weights.bin` layout (`/root/reference/experiment/mnist/model.json`,
SURVEY §5.1) maps onto these modules without any transpose.

Forward/backward of Dense, Conv2D, MaxPooling2D go through `distriflow_b200.ops`,
which runs the hand-written sm_100a kernels on CUDA tensors and a PyTorch
restatement on CPU tensors.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops

__all__ = ["Dense", "Conv2D", "MaxPooling2D", "AveragePooling2D", "GlobalAveragePooling2D", "Flatten",
           "Dropout", "Activation", "BatchNormalization", "LayerNorm", "apply_activation"]


def apply_activation(x: torch.Tensor, name: Optional[str]) -> torch.Tensor:
    if name in (None, "linear"):
        return x
    if name == "relu":
        return F.relu(x)
    if name == "softmax":
        return F.softmax(x.float(), dim=-1).to(x.dtype)
    if name == "sigmoid":
        return torch.sigmoid(x)
    if name == "tanh":
        return torch.tanh(x)
    if name == "gelu":
        return F.gelu(x, approximate="tanh")
    raise ValueError(f"unsupported activation {name}")


def _glorot_uniform_(t: torch.Tensor, fan_in: int, fan_out: int) -> None:
    # VarianceScaling(scale=1, mode=fan_avg, distribution=uniform) as in model.json
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        t.uniform_(-limit, limit)


class Dense(nn.Module):
    def __init__(self, in_features: int, units: int, activation: Optional[str] = None, use_bias: bool = True,
                 name: str = "dense"):
        super().__init__()
        self.in_features, self.units, self.activation, self.layer_name = in_features, units, activation, name
        self.kernel = nn.Parameter(torch.empty(in_features, units))
        self.bias = nn.Parameter(torch.zeros(units)) if use_bias else None
        _glorot_uniform_(self.kernel, in_features, units)

    def forward(self, x):
        fused = self.activation if self.activation in ("relu", "gelu") else None
        y = ops.linear(x, self.kernel, self.bias, fused)
        return y if fused == self.activation else apply_activation(y, self.activation)

    def extra_repr(self):
        return f"{self.in_features}->{self.units}, activation={self.activation}"


class Conv2D(nn.Module):
    """NHWC conv with HWIO filters; `padding` in {'valid','same'}."""

    def __init__(self, in_channels: int, filters: int, kernel_size, strides=1, padding: str = "valid",
                 activation: Optional[str] = None, use_bias: bool = True, name: str = "conv2d"):
        super().__init__()
        ks = (kernel_size, kernel_size) if isinstance(kernel_size, int) else tuple(kernel_size)
        st = (strides, strides) if isinstance(strides, int) else tuple(strides)
        self.in_channels, self.filters, self.kernel_size, self.strides = in_channels, filters, ks, st
        self.padding, self.activation, self.layer_name = padding, activation, name
        self.kernel = nn.Parameter(torch.empty(ks[0], ks[1], in_channels, filters))
        self.bias = nn.Parameter(torch.zeros(filters)) if use_bias else None
        rf = ks[0] * ks[1]
        _glorot_uniform_(self.kernel, in_channels * rf, filters * rf)

    def forward(self, x):
        fused = self.activation if self.activation == "relu" else None
        y = ops.conv2d_nhwc(x, self.kernel, self.bias, self.strides, self.padding, fused)
        return y if fused == self.activation else apply_activation(y, self.activation)


class MaxPooling2D(nn.Module):
    def __init__(self, pool_size=2, strides=None, padding: str = "valid", name: str = "max_pooling2d"):
        super().__init__()
        ps = (pool_size, pool_size) if isinstance(pool_size, int) else tuple(pool_size)
        st = ps if strides is None else ((strides, strides) if isinstance(strides, int) else tuple(strides))
        self.pool_size, self.strides, self.padding, self.layer_name = ps, st, padding, name

    def forward(self, x):
        return ops.maxpool2d_nhwc(x, self.pool_size, self.strides, self.padding)


class AveragePooling2D(nn.Module):
    def __init__(self, pool_size=2, strides=None, name: str = "average_pooling2d"):
        super().__init__()
        ps = (pool_size, pool_size) if isinstance(pool_size, int) else tuple(pool_size)
        self.pool_size, self.strides, self.layer_name = ps, (ps if strides is None else tuple(strides)), name

    def forward(self, x):
        y = F.avg_pool2d(x.permute(0, 3, 1, 2), self.pool_size, self.strides)
        return y.permute(0, 2, 3, 1).contiguous()


class GlobalAveragePooling2D(nn.Module):
    def __init__(self, name: str = "global_average_pooling2d"):
        super().__init__()
        self.layer_name = name

    def forward(self, x):
        return x.float().mean(dim=(1, 2)).to(x.dtype)


class Flatten(nn.Module):
    def __init__(self, name: str = "flatten"):
        super().__init__()
        self.layer_name = name

    def forward(self, x):
        return x.reshape(x.shape[0], -1)


class Dropout(nn.Module):
    def __init__(self, rate: float, name: str = "dropout"):
        super().__init__()
        self.rate, self.layer_name = rate, name

    def forward(self, x):
        return F.dropout(x, self.rate, self.training)


class Activation(nn.Module):
    def __init__(self, activation: str, name: str = "activation"):
        super().__init__()
        self.activation, self.layer_name = activation, name

    def forward(self, x):
        return apply_activation(x, self.activation)


class BatchNormalization(nn.Module):
    """Channels-last batch norm (axis=-1); weight names follow Keras: gamma, beta, moving_mean, moving_variance."""

    def __init__(self, channels: int, momentum: float = 0.99, epsilon: float = 1e-3, name: str = "batch_normalization"):
        super().__init__()
        self.channels, self.momentum, self.epsilon, self.layer_name = channels, momentum, epsilon, name
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))
        self.register_buffer("moving_mean", torch.zeros(channels))
        self.register_buffer("moving_variance", torch.ones(channels))

    def forward(self, x, residual=None, activation=None):
        """`activation(BN(x) + residual)`: the add and the ReLU of a ResNet block ride on the normalisation kernel."""
        return ops.batch_norm_nhwc(x, self.gamma, self.beta, self.moving_mean, self.moving_variance, self.training,
                                   1.0 - self.momentum, self.epsilon, activation, residual)


class LayerNorm(nn.Module):
    def __init__(self, dim: int, epsilon: float = 1e-5, name: str = "layer_normalization"):
        super().__init__()
        self.dim, self.epsilon, self.layer_name = dim, epsilon, name
        self.gamma = nn.Parameter(torch.ones(dim))
        self.beta = nn.Parameter(torch.zeros(dim))

    def forward(self, x):
        return ops.layer_norm(x, self.gamma, self.beta, self.epsilon)
