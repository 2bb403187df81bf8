"""Keras-style `Sequential` topology <-> torch module.

The reference stores models as tf.js LayersModel artifacts whose
`modelTopology` is a Keras config (in-repo example:
`/root/reference/experiment/mnist/model.json`; builder form:
`/root/reference/experiment/mnist/mnist_server.ts:16-22`).  This module builds
a `SequentialModel` from such a config and emits the config back from a model,
so checkpoints round-trip.  Both shapes of topology are accepted:
the converter style (`{"model_config": {"class_name": "Sequential", "config": [...]}, "keras_version", ...}`)
and the tfjs-saved style (`{"class_name": "Sequential", "config": {"layers": [...]}}`).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from . import layers as L

__all__ = ["SequentialModel", "model_from_topology", "topology_from_model", "sequential"]


class SequentialModel(nn.Module):
    """A list of layers + the batch-less input shape (NHWC for images)."""

    def __init__(self, layer_list: Sequence[nn.Module], input_shape: Sequence[int], name: str = "sequential_1"):
        super().__init__()
        self.layers = nn.ModuleList(layer_list)
        self.input_shape = list(input_shape)
        self.model_name = name
        self.output_shape = self._infer_output_shape()

    def forward(self, x):
        if x.dim() == len(self.input_shape):        # single example
            x = x.unsqueeze(0)
        for layer in self.layers:
            x = layer(x)
        return x

    def _infer_output_shape(self) -> List[int]:
        was = self.training
        self.eval()
        with torch.no_grad():
            p = next(self.parameters(), None)
            dt = p.dtype if p is not None else torch.float32
            y = self.forward(torch.zeros(1, *self.input_shape, dtype=dt))
        self.train(was)
        return list(y.shape[1:])

    def first_dense_kernel(self):
        for layer in self.layers:
            if isinstance(layer, L.Dense):
                return layer.kernel
            if isinstance(layer, L.Conv2D):
                return None
        return None

    # weight naming follows tfjs: "<layer_name>/kernel|bias|gamma|beta|moving_mean|moving_variance"
    def named_weights(self) -> List[Tuple[str, torch.Tensor]]:
        out = []
        for layer in self.layers:
            lname = getattr(layer, "layer_name", layer.__class__.__name__.lower())
            for pname, p in list(layer.named_parameters(recurse=False)) + list(layer.named_buffers(recurse=False)):
                out.append((f"{lname}/{pname}", p))
        return out


def sequential(layer_list: Sequence[nn.Module], input_shape: Sequence[int], name: str = "sequential_1") -> SequentialModel:
    return SequentialModel(layer_list, input_shape, name)


def is_sequential_topology(topology: Dict[str, Any]) -> bool:
    """True when `model_from_topology` can rebuild the module from this JSON (Keras `Sequential`)."""
    t = topology.get("model_config", topology)
    return isinstance(t, dict) and t.get("class_name") == "Sequential"


def _layer_entries(topology: Dict[str, Any]) -> Tuple[List[Dict[str, Any]], str]:
    t = topology
    if "model_config" in t:
        t = t["model_config"]
    if t.get("class_name") != "Sequential":
        raise ValueError(f"only Sequential topologies are supported, got {t.get('class_name')}")
    cfg = t["config"]
    if isinstance(cfg, dict):
        return list(cfg["layers"]), cfg.get("name", "sequential_1")
    return list(cfg), "sequential_1"


def model_from_topology(topology: Dict[str, Any]) -> SequentialModel:
    entries, name = _layer_entries(topology)
    shape: Optional[List[int]] = None          # running batch-less shape
    input_shape: Optional[List[int]] = None
    mods: List[nn.Module] = []
    for e in entries:
        cls, c = e["class_name"], e["config"]
        if shape is None:
            bis = c.get("batch_input_shape") or c.get("batchInputShape")
            if cls == "InputLayer":
                input_shape = shape = list(bis[1:])
                continue
            if bis is None:
                raise ValueError("first layer needs batch_input_shape")
            input_shape = shape = list(bis[1:])
        lname = c.get("name", cls.lower())
        act = c.get("activation")
        if cls == "Dense":
            if len(shape) != 1:
                raise ValueError(f"Dense after shape {shape}: add Flatten")
            mods.append(L.Dense(shape[0], int(c["units"]), act, c.get("use_bias", c.get("useBias", True)), lname))
            shape = [int(c["units"])]
        elif cls == "Conv2D":
            ks = c.get("kernel_size") or c.get("kernelSize")
            st = c.get("strides", [1, 1])
            pad = c.get("padding", "valid")
            mods.append(L.Conv2D(shape[2], int(c["filters"]), tuple(ks), tuple(st), pad, act,
                                 c.get("use_bias", c.get("useBias", True)), lname))
            shape = _conv_out(shape, ks, st, pad) + [int(c["filters"])]
        elif cls == "MaxPooling2D":
            ps = c.get("pool_size") or c.get("poolSize") or [2, 2]
            st = c.get("strides") or ps
            pad = c.get("padding", "valid")
            mods.append(L.MaxPooling2D(tuple(ps), tuple(st), pad, lname))
            shape = _conv_out(shape, ps, st, pad) + [shape[2]]
        elif cls == "AveragePooling2D":
            ps = c.get("pool_size") or c.get("poolSize") or [2, 2]
            st = c.get("strides") or ps
            mods.append(L.AveragePooling2D(tuple(ps), tuple(st), lname))
            shape = _conv_out(shape, ps, st, "valid") + [shape[2]]
        elif cls == "GlobalAveragePooling2D":
            mods.append(L.GlobalAveragePooling2D(lname))
            shape = [shape[2]]
        elif cls == "Flatten":
            mods.append(L.Flatten(lname))
            n = 1
            for d in shape:
                n *= d
            shape = [n]
        elif cls == "Dropout":
            mods.append(L.Dropout(float(c["rate"]), lname))
        elif cls == "Activation":
            mods.append(L.Activation(c["activation"], lname))
        elif cls == "BatchNormalization":
            mods.append(L.BatchNormalization(shape[-1], c.get("momentum", 0.99), c.get("epsilon", 1e-3), lname))
        elif cls == "ReLU":
            mods.append(L.Activation("relu", lname))
        elif cls == "Softmax":
            mods.append(L.Activation("softmax", lname))
        else:
            raise ValueError(f"unsupported layer class {cls}")
    return SequentialModel(mods, input_shape, name)


def _conv_out(shape, ks, st, pad) -> List[int]:
    h, w = shape[0], shape[1]
    if pad == "same":
        return [-(-h // st[0]), -(-w // st[1])]
    return [(h - ks[0]) // st[0] + 1, (w - ks[1]) // st[1] + 1]


_INIT = {"class_name": "VarianceScaling",
         "config": {"distribution": "uniform", "scale": 1.0, "seed": None, "mode": "fan_avg"}}


def topology_from_model(model: SequentialModel) -> Dict[str, Any]:
    entries: List[Dict[str, Any]] = []
    for i, layer in enumerate(model.layers):
        name = getattr(layer, "layer_name", f"layer_{i}")
        if isinstance(layer, L.Dense):
            c = {"name": name, "units": layer.units, "activation": layer.activation or "linear",
                 "use_bias": layer.bias is not None, "trainable": True, "kernel_initializer": _INIT,
                 "bias_initializer": {"class_name": "Zeros", "config": {}}}
            cls = "Dense"
        elif isinstance(layer, L.Conv2D):
            c = {"name": name, "filters": layer.filters, "kernel_size": list(layer.kernel_size),
                 "strides": list(layer.strides), "padding": layer.padding, "data_format": "channels_last",
                 "dilation_rate": [1, 1], "activation": layer.activation or "linear",
                 "use_bias": layer.bias is not None, "trainable": True, "kernel_initializer": _INIT,
                 "bias_initializer": {"class_name": "Zeros", "config": {}}}
            cls = "Conv2D"
        elif isinstance(layer, L.MaxPooling2D):
            c = {"name": name, "pool_size": list(layer.pool_size), "strides": list(layer.strides),
                 "padding": layer.padding, "data_format": "channels_last", "trainable": True}
            cls = "MaxPooling2D"
        elif isinstance(layer, L.AveragePooling2D):
            c = {"name": name, "pool_size": list(layer.pool_size), "strides": list(layer.strides),
                 "padding": "valid", "data_format": "channels_last", "trainable": True}
            cls = "AveragePooling2D"
        elif isinstance(layer, L.GlobalAveragePooling2D):
            c, cls = {"name": name, "data_format": "channels_last", "trainable": True}, "GlobalAveragePooling2D"
        elif isinstance(layer, L.Flatten):
            c, cls = {"name": name, "trainable": True}, "Flatten"
        elif isinstance(layer, L.Dropout):
            c, cls = {"name": name, "rate": layer.rate, "noise_shape": None, "seed": None, "trainable": True}, "Dropout"
        elif isinstance(layer, L.Activation):
            c, cls = {"name": name, "activation": layer.activation, "trainable": True}, "Activation"
        elif isinstance(layer, L.BatchNormalization):
            c = {"name": name, "axis": -1, "momentum": layer.momentum, "epsilon": layer.epsilon, "center": True,
                 "scale": True, "trainable": True}
            cls = "BatchNormalization"
        else:
            raise ValueError(f"cannot express {layer.__class__.__name__} as a Keras layer config")
        if i == 0:
            c["batch_input_shape"] = [None, *model.input_shape]
            c["dtype"] = "float32"
        entries.append({"class_name": cls, "config": c})
    return {"class_name": "Sequential", "config": {"name": model.model_name, "layers": entries},
            "keras_version": "tfjs-layers 1.2.9", "backend": "tensor_flow.js"}
