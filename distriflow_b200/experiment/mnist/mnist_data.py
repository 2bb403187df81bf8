"""MNIST IDX loader -> DistriDataset.

Parity: `/root/reference/experiment/mnist/mnist_data.ts:1-75` — reads `train-images-idx3-ubyte` / `train-labels-idx1-ubyte`
(and the t10k pair) from a `data/` directory, checks the magics 0x803 / 0x801, images become float32 `[N,28,28,1]` with
RAW 0..255 values (no normalisation), labels become one-hot over 10 classes, and the training split is wrapped
in a dataset dispatcher with the default config (batch 32, 5 epochs).

Reference quirk not reproduced by default: the loader byte-swaps the WHOLE file with `swap32()` before slicing out the
pixel bytes (`mnist_data.ts:22,41`), which also permutes every group of four pixels; `swap32_quirk=True` reproduces it.
The data files are not part of either repository (`.gitignore`); without them `synthetic=True` builds MNIST-shaped
class-dependent blobs so the experiment still runs offline.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ...config import DEFAULT_DATASET_HYPERPARAMS
from ...dataset import DistriDataset
from ...oracle import synthetic_mnist

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
TRAIN = ("train-images-idx3-ubyte", "train-labels-idx1-ubyte")
TEST = ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte")


def load_mnist_format(imgs_path: str, labels_path: str, swap32_quirk: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    raw_i = open(imgs_path, "rb").read()
    raw_l = open(labels_path, "rb").read()
    magic_i, n, rows, cols = struct.unpack(">IIII", raw_i[:16])
    magic_l, n_l = struct.unpack(">II", raw_l[:8])
    if magic_i != 0x00000803:
        raise ValueError(f"Training images file has invalid magic number 0x00000803 !== {magic_i:x}")
    if magic_l != 0x00000801:
        raise ValueError(f"Training labels file has invalid magic number 0x00000801 !== {magic_l:x}")
    if n_l != n:
        raise ValueError(f"{n} images but {n_l} labels")
    px = np.frombuffer(raw_i, dtype=np.uint8, count=n * rows * cols, offset=16)
    lab = np.frombuffer(raw_l, dtype=np.uint8, count=n, offset=8)
    if swap32_quirk:
        px = px[: (px.size // 4) * 4].reshape(-1, 4)[:, ::-1].reshape(-1)
        lab = np.concatenate([lab[: (lab.size // 4) * 4].reshape(-1, 4)[:, ::-1].reshape(-1), lab[(lab.size // 4) * 4:]])
    imgs = torch.from_numpy(px.astype(np.float32).reshape(n, rows, cols, 1))
    labels = torch.from_numpy(lab.astype(np.int32))
    return imgs, labels


def load_mnist(data_dir: str = DATA_DIR, swap32_quirk: bool = False) -> Dict[str, Tuple[torch.Tensor, torch.Tensor]]:
    return {"train": load_mnist_format(os.path.join(data_dir, TRAIN[0]), os.path.join(data_dir, TRAIN[1]), swap32_quirk),
            "val": load_mnist_format(os.path.join(data_dir, TEST[0]), os.path.join(data_dir, TEST[1]), swap32_quirk)}


def load_dataset(data_dir: str = DATA_DIR, config: Optional[dict] = None, synthetic: Optional[bool] = None,
                 n_synthetic: int = 60000, seed: int = 0) -> DistriDataset:
    have = all(os.path.exists(os.path.join(data_dir, f)) for f in TRAIN)
    if synthetic is None:
        synthetic = not have
    if synthetic:
        x, y = synthetic_mnist(n_synthetic, seed=seed)
        imgs, onehot = torch.from_numpy(x), torch.from_numpy(y)
    else:
        imgs, labels = load_mnist_format(os.path.join(data_dir, TRAIN[0]), os.path.join(data_dir, TRAIN[1]))
        onehot = torch.nn.functional.one_hot(labels.long(), 10).to(torch.int32)          # tf.oneHot -> int32 (mnist_data.ts:70)
    return DistriDataset(imgs, onehot, config or dict(DEFAULT_DATASET_HYPERPARAMS))
