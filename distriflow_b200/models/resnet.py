"""ResNet-20 (CIFAR-10 shape, BASELINE.json config 3) and ResNet-50 v1.5 (224x224, config 4), channels-last.

Not in the reference (it ships only the MNIST models).  Built from this package's own layers: Conv2D with HWIO
filters runs the hand-written implicit-GEMM kernels (tcgen05 for bf16 with C_in % 64 == 0, CUDA-core otherwise),
the classifier head is the Dense GEMM, the loss is the fused softmax-CE, and BatchNorm with the residual add and the
ReLU of a block is one fused op (`ops.batch_norm_nhwc`, `csrc/batchnorm.cu`).  Global average pooling stays on PyTorch.
"""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.nn as nn

from .. import ops
from . import layers as L


class ConvBN(nn.Module):
    def __init__(self, cin: int, cout: int, k: int, stride: int = 1, relu: bool = True, name: str = "conv"):
        super().__init__()
        self.conv = L.Conv2D(cin, cout, k, stride, "same", None, use_bias=False, name=name)
        self.bn = L.BatchNormalization(cout, momentum=0.9, epsilon=1e-5, name=name + "_bn")
        self.relu = relu
        nn.init.kaiming_normal_(self.conv.kernel.data.permute(3, 2, 0, 1), mode="fan_out", nonlinearity="relu")

    def forward(self, x, residual=None, relu=None):
        """`relu(bn(conv(x)) + residual)`; the add and the ReLU are fused into the BatchNorm kernels."""
        relu = self.relu if relu is None else relu
        return self.bn(self.conv(x), residual, "relu" if relu else None)


class BasicBlock(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int, name: str):
        super().__init__()
        self.c1 = ConvBN(cin, cout, 3, stride, True, name + "/c1")
        self.c2 = ConvBN(cout, cout, 3, 1, False, name + "/c2")
        self.short = None if (stride == 1 and cin == cout) else ConvBN(cin, cout, 1, stride, False, name + "/short")

    def forward(self, x):
        return self.c2(self.c1(x), x if self.short is None else self.short(x), True)


class Bottleneck(nn.Module):
    def __init__(self, cin: int, mid: int, stride: int, name: str):
        super().__init__()
        cout = mid * 4
        self.c1 = ConvBN(cin, mid, 1, 1, True, name + "/c1")
        self.c2 = ConvBN(mid, mid, 3, stride, True, name + "/c2")          # v1.5: stride on the 3x3
        self.c3 = ConvBN(mid, cout, 1, 1, False, name + "/c3")
        self.short = None if (stride == 1 and cin == cout) else ConvBN(cin, cout, 1, stride, False, name + "/short")

    def forward(self, x):
        return self.c3(self.c2(self.c1(x)), x if self.short is None else self.short(x), True)


class ResNet(nn.Module):
    def __init__(self, stem: nn.Module, stages: Sequence[nn.Module], features: int, classes: int, input_shape: List[int]):
        super().__init__()
        self.stem = stem
        self.stages = nn.ModuleList(stages)
        self.pool = L.GlobalAveragePooling2D()
        self.fc = L.Dense(features, classes, name="fc")
        self.input_shape = input_shape
        self.output_shape = [classes]

    def forward(self, x):
        p = next(self.parameters())
        if x.dtype != p.dtype:
            x = x.to(p.dtype)
        x = self.stem(x)
        for s in self.stages:
            x = s(x)
        return self.fc(self.pool(x))


def resnet20(classes: int = 10) -> ResNet:
    stem = ConvBN(3, 16, 3, 1, True, "stem")
    stages, cin = [], 16
    for si, c in enumerate((16, 32, 64)):
        blocks = []
        for bi in range(3):
            blocks.append(BasicBlock(cin, c, 2 if (bi == 0 and si > 0) else 1, f"s{si}b{bi}"))
            cin = c
        stages.append(nn.Sequential(*blocks))
    return ResNet(stem, stages, 64, classes, [32, 32, 3])


class _Stem50(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = ConvBN(3, 64, 7, 2, True, "stem")
        self.pool = L.MaxPooling2D(3, 2, "same", "stem_pool")

    def forward(self, x):
        return self.pool(self.conv(x))


def resnet50(classes: int = 1000) -> ResNet:
    stages, cin = [], 64
    for si, (mid, n) in enumerate(((64, 3), (128, 4), (256, 6), (512, 3))):
        blocks = []
        for bi in range(n):
            blocks.append(Bottleneck(cin, mid, 2 if (bi == 0 and si > 0) else 1, f"s{si}b{bi}"))
            cin = mid * 4
        stages.append(nn.Sequential(*blocks))
    return ResNet(_Stem50(), stages, 2048, classes, [224, 224, 3])


def resnet_loss(model: ResNet, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    return ops.softmax_cross_entropy(model(x), y)
