"""The two MNIST models of the reference experiment.

`mnist_mlp`: `Flatten[28,28,1] -> Dense(10, relu) -> Dense(10, softmax)`, 7,960 parameters
(`/root/reference/experiment/mnist/mnist_server.ts:16-22`, same in `mnist_client.ts:15-21`).
`mnist_cnn`: the Keras-converted CNN of `/root/reference/experiment/mnist/model.json`
(Conv32-ReLU-Conv32-ReLU-MaxPool-Dropout-Flatten-Dense128-ReLU-Dropout-Dense(classes)-softmax);
the JSON in the repo has 5 output units, 600,165 parameters.
"""
from __future__ import annotations

from . import layers as L
from .topology import SequentialModel


def mnist_mlp(softmax_output: bool = True) -> SequentialModel:
    return SequentialModel([
        L.Flatten("flatten_1"),
        L.Dense(784, 10, "relu", name="dense_1"),
        L.Dense(10, 10, "softmax" if softmax_output else None, name="dense_2"),
    ], [28, 28, 1], "mnist_dense")


def mnist_cnn(classes: int = 5, softmax_output: bool = True, dropout: bool = True) -> SequentialModel:
    layers = [
        L.Conv2D(1, 32, 3, name="conv2d_1"), L.Activation("relu", "activation_1"),
        L.Conv2D(32, 32, 3, name="conv2d_2"), L.Activation("relu", "activation_2"),
        L.MaxPooling2D(2, name="max_pooling2d_1"), L.Dropout(0.25 if dropout else 0.0, "dropout_1"),
        L.Flatten("flatten_1"),
        L.Dense(4608, 128, name="dense_1"), L.Activation("relu", "activation_3"),
        L.Dropout(0.5 if dropout else 0.0, "dropout_2"),
        L.Dense(128, classes, name="dense_2"),
    ]
    if softmax_output:
        layers.append(L.Activation("softmax", "activation_4"))
    return SequentialModel(layers, [28, 28, 1], "mnist_cnn")
