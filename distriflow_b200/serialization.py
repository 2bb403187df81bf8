"""Tensor (de)serialisation, update stacking and the flat tensor file format.

Parity targets in the reference:
  * `SerializedVariable {dtype, shape, data}` and `serializeVar(s)` /
    `deserializeVar(s)` / `stackSerialized` / `serializedToArray`
    (`/root/reference/src/common/utils.ts:5-101`),
  * `flatSerialize` / `flatDeserialize` (`/root/reference/src/server/models.ts:220-267`).

The wire form is used by the host-side transports (in-process bus, TCP) and by
the checkpoint writer.  On the GPU data path nothing is serialised: gradients
and weights move through peer-mapped buffers (see `parallel/`).
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Any, Dict, Iterable, List, Sequence, Tuple

import numpy as np
import torch

__all__ = [
    "SerializedVariable", "serialize_var", "serialize_vars", "deserialize_var",
    "deserialize_vars", "stack_serialized", "serialized_to_array",
    "flat_serialize", "flat_deserialize", "DTYPES",
]

# wire dtype name -> (numpy dtype, torch dtype).  float32/int32/bool are the
# reference's set (utils.ts:13-17); the 16-bit floats are extensions used by
# the bf16 configs.  bfloat16 has no numpy dtype: it travels as raw uint16.
DTYPES: Dict[str, Tuple[Any, torch.dtype]] = {
    "float32": (np.float32, torch.float32),
    "int32": (np.int32, torch.int32),
    "bool": (np.uint8, torch.bool),
    "float16": (np.float16, torch.float16),
    "bfloat16": (np.uint16, torch.bfloat16),
    "int64": (np.int64, torch.int64),
}
_TORCH_TO_NAME = {v[1]: k for k, v in DTYPES.items()}


@dataclass
class SerializedVariable:
    dtype: str
    shape: Tuple[int, ...]
    data: bytes

    def nbytes(self) -> int:
        return len(self.data)

    def to_json(self) -> Dict[str, Any]:
        return {"dtype": self.dtype, "shape": list(self.shape)}


def _as_tensor(v: Any) -> torch.Tensor:
    if isinstance(v, torch.Tensor):
        return v
    if hasattr(v, "read"):           # LayerVariable-like (utils.ts:39-51)
        return _as_tensor(v.read())
    if hasattr(v, "data") and isinstance(getattr(v, "data"), torch.Tensor):
        return v.data
    arr = np.asarray(v)
    if arr.dtype == np.float64:
        arr = arr.astype(np.float32)
    if arr.dtype == np.bool_:
        return torch.from_numpy(arr.astype(np.uint8)).to(torch.bool)
    return torch.from_numpy(arr)


def serialize_var(variable: Any) -> SerializedVariable:
    """Device->host readback + raw little-endian copy (utils.ts:32-37)."""
    t = _as_tensor(variable).detach()
    if t.dtype == torch.float64:
        t = t.to(torch.float32)
    if t.dtype not in _TORCH_TO_NAME:
        raise TypeError(f"unsupported dtype {t.dtype}")
    name = _TORCH_TO_NAME[t.dtype]
    t = t.contiguous().cpu()
    if t.dtype == torch.bfloat16:
        raw = t.view(torch.uint16).numpy().tobytes()
    elif t.dtype == torch.bool:
        raw = t.to(torch.uint8).numpy().tobytes()
    else:
        raw = t.numpy().tobytes()
    return SerializedVariable(name, tuple(int(d) for d in t.shape), raw)


def serialize_vars(variables: Iterable[Any]) -> List[SerializedVariable]:
    return [serialize_var(v) for v in variables]


def serialized_to_array(s: SerializedVariable) -> np.ndarray:
    """Typed view over the payload; accepts bytes/bytearray/memoryview (utils.ts:86-101)."""
    if s.dtype not in DTYPES:
        raise TypeError(f"unsupported dtype {s.dtype}")
    np_dtype = DTYPES[s.dtype][0]
    n = int(np.prod(s.shape)) if len(s.shape) else 1
    arr = np.frombuffer(bytes(s.data) if not isinstance(s.data, (bytes, bytearray)) else s.data,
                        dtype=np_dtype)
    if arr.size != n:
        raise ValueError(f"payload has {arr.size} elements, shape {s.shape} needs {n}")
    return arr


def deserialize_var(s: SerializedVariable, device: Any = None) -> torch.Tensor:
    arr = serialized_to_array(s).copy()
    torch_dtype = DTYPES[s.dtype][1]
    t = torch.from_numpy(arr)
    if torch_dtype == torch.bfloat16:
        t = t.view(torch.bfloat16)
    elif torch_dtype == torch.bool:
        t = t.to(torch.bool)
    t = t.reshape(s.shape)
    return t.to(device) if device is not None else t


def deserialize_vars(vs: Iterable[SerializedVariable], device: Any = None) -> List[torch.Tensor]:
    return [deserialize_var(v, device) for v in vs]


def stack_serialized(updates: Sequence[Sequence[SerializedVariable]]) -> List[SerializedVariable]:
    """K updates x W weights -> W variables of shape [K, ...] (utils.ts:53-75).

    Byte concatenation per weight; this is the host-side stand-in for the
    device reduce-scatter (the GPU path never stacks)."""
    if not updates:
        return []
    k = len(updates)
    out: List[SerializedVariable] = []
    for w, first in enumerate(updates[0]):
        for u in updates:
            if u[w].dtype != first.dtype or tuple(u[w].shape) != tuple(first.shape):
                raise ValueError(f"update mismatch for weight {w}")
        out.append(SerializedVariable(first.dtype, (k, *first.shape),
                                      b"".join(bytes(u[w].data) for u in updates)))
    return out


# ----------------------------------------------------------------------------------------------
# flat format: one data blob + JSON {meta:[{shape,dtype}], byteOffsets:[...]} (server/models.ts:220-267)

def flat_serialize(tensors: Sequence[Any]) -> Tuple[bytes, Dict[str, Any]]:
    metas, offsets, chunks, off = [], [], [], 0
    for t in tensors:
        s = serialize_var(t)
        metas.append({"shape": list(s.shape), "dtype": s.dtype})
        offsets.append(off)
        chunks.append(s.data)
        off += len(s.data)
    return b"".join(chunks), {"meta": metas, "byteOffsets": offsets}


def flat_deserialize(data: bytes, meta: Dict[str, Any], device: Any = None) -> List[torch.Tensor]:
    out = []
    offs = list(meta["byteOffsets"]) + [len(data)]
    for i, m in enumerate(meta["meta"]):
        np_dtype = np.dtype(DTYPES[m["dtype"]][0])
        n = int(np.prod(m["shape"])) if len(m["shape"]) else 1
        beg = offs[i]
        end = beg + n * np_dtype.itemsize
        if end > offs[i + 1] and i + 1 < len(offs):
            raise ValueError("flat blob truncated")
        out.append(deserialize_var(SerializedVariable(m["dtype"], tuple(m["shape"]), data[beg:end]), device))
    return out


def flat_meta_dumps(meta: Dict[str, Any]) -> str:
    return json.dumps(meta)
