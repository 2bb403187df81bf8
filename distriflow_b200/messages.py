"""Wire-protocol message types shared by servers and workers.

Same two events and the same message shapes as the reference
(`/root/reference/src/common/utils.ts:103-155`): a `Download` carries
`{model:{version, vars}, hyperparams, data?}`, an `Upload` carries
`{clientId, gradients:{version, vars}?, batch?, metrics?}`.

They are used by the host transports (in-process, TCP).  Between GPU ranks of
one box the same information travels as device-resident state: versions are
device counters, `vars` are peer-mapped shards, `batch` is a work-queue ticket.
"""
from __future__ import annotations

import json
import struct
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional

from .serialization import SerializedVariable

__all__ = ["Events", "ModelMsg", "GradientMsg", "DataMsg", "Batch", "UploadMsg",
           "DownloadMsg", "VersionCallback", "UploadCallback", "PreprocessCallback",
           "encode", "decode", "WireFormatError", "MAX_FRAME_BYTES"]


class Events:
    """Event names (utils.ts:115-118) plus the connection lifecycle events."""
    Download = "downloadVars"
    Upload = "uploadVars"
    Connect = "connection"
    Disconnect = "disconnect"


@dataclass
class ModelMsg:           # utils.ts:120-123
    version: str
    vars: List[SerializedVariable]


@dataclass
class GradientMsg:        # utils.ts:125-128
    version: str
    vars: List[SerializedVariable]


@dataclass
class DataMsg:            # utils.ts:130-135 (+ uuid ids from README.md:33-45)
    batch: int
    epoch: int
    x: SerializedVariable
    y: SerializedVariable
    batchId: Optional[str] = None
    epochId: Optional[str] = None


@dataclass
class Batch:              # utils.ts:137-142
    batch: int
    epoch: int
    x: Any
    y: Any
    batchId: Optional[str] = None
    epochId: Optional[str] = None
    done: bool = False


@dataclass
class UploadMsg:          # utils.ts:144-149
    clientId: str
    gradients: Optional[GradientMsg] = None
    batch: Optional[int] = None
    metrics: Optional[List[float]] = None
    epoch: Optional[int] = None


@dataclass
class DownloadMsg:        # utils.ts:151-155
    model: ModelMsg
    hyperparams: Dict[str, Any] = field(default_factory=dict)
    data: Optional[DataMsg] = None


VersionCallback = Callable[[Optional[str], str], None]       # (oldVersion, newVersion)
UploadCallback = Callable[[Optional[UploadMsg]], None]
PreprocessCallback = Callable[[Batch], Batch]


# ----------------------------------------------------------------------------------------------
# Data-only wire codec.  The reference protocol is built for untrusted peers (browsers over socket.io:
# JSON + ArrayBuffer attachments), so a frame must never be able to run code on the receiver.  A frame is
#     u32 header length | JSON header {"t": tree, "b": [blob lengths]} | blobs back to back
# where `tree` holds only null / bool / number / string / list / dict and tagged nodes for tuples, byte blobs,
# `SerializedVariable`s and the message dataclasses above (by name, from a fixed table).
MAX_FRAME_BYTES = 1 << 31          # the transport refuses anything larger before allocating
_MSG_TYPES = {c.__name__: c for c in (ModelMsg, GradientMsg, DataMsg, Batch, UploadMsg, DownloadMsg)}


class WireFormatError(ValueError):
    pass


def _pack(obj: Any, blobs: List[bytes]) -> Any:
    if obj is None or isinstance(obj, (bool, int, float, str)):
        return obj
    if isinstance(obj, (bytes, bytearray, memoryview)):
        blobs.append(bytes(obj))
        return {"$": "b", "i": len(blobs) - 1}
    if isinstance(obj, SerializedVariable):
        blobs.append(bytes(obj.data))
        return {"$": "v", "d": obj.dtype, "s": list(obj.shape), "i": len(blobs) - 1}
    if isinstance(obj, tuple):
        return {"$": "t", "v": [_pack(o, blobs) for o in obj]}
    if isinstance(obj, list):
        return [_pack(o, blobs) for o in obj]
    if isinstance(obj, dict):
        for k in obj:
            if not isinstance(k, str):
                raise WireFormatError(f"dict keys must be strings on the wire, got {type(k).__name__}")
        return {"$": "d", "v": {k: _pack(v, blobs) for k, v in obj.items()}}
    name = type(obj).__name__
    if _MSG_TYPES.get(name) is type(obj):
        return {"$": "m", "n": name, "v": {k: _pack(getattr(obj, k), blobs) for k in obj.__dataclass_fields__}}
    if hasattr(obj, "item") and getattr(obj, "ndim", 1) == 0:          # numpy / torch scalars
        return _pack(obj.item(), blobs)
    raise WireFormatError(f"{type(obj).__name__} cannot travel on the wire (data-only protocol)")


def _unpack(node: Any, blobs: List[bytes]) -> Any:
    if isinstance(node, list):
        return [_unpack(n, blobs) for n in node]
    if not isinstance(node, dict):
        return node
    tag = node.get("$")
    try:
        if tag == "b":
            return blobs[node["i"]]
        if tag == "v":
            return SerializedVariable(str(node["d"]), tuple(int(x) for x in node["s"]), blobs[node["i"]])
        if tag == "t":
            return tuple(_unpack(n, blobs) for n in node["v"])
        if tag == "d":
            return {str(k): _unpack(v, blobs) for k, v in node["v"].items()}
        if tag == "m":
            cls = _MSG_TYPES[node["n"]]
            fields = {k: _unpack(v, blobs) for k, v in node["v"].items() if k in cls.__dataclass_fields__}
            return cls(**fields)
    except (KeyError, IndexError, TypeError) as e:
        raise WireFormatError(f"malformed frame: {e!r}") from None
    raise WireFormatError(f"unknown node tag {tag!r}")


def encode(obj: Any) -> bytes:
    """Binary framing payload for the TCP transport (data only: nothing in a frame is executable)."""
    blobs: List[bytes] = []
    tree = _pack(obj, blobs)
    head = json.dumps({"t": tree, "b": [len(b) for b in blobs]}, separators=(",", ":")).encode()
    return b"".join([struct.pack("<I", len(head)), head, *blobs])


def decode(buf: bytes) -> Any:
    if len(buf) < 4:
        raise WireFormatError("short frame")
    (n,) = struct.unpack_from("<I", buf, 0)
    if 4 + n > len(buf):
        raise WireFormatError("header runs past the frame")
    try:
        head = json.loads(bytes(buf[4:4 + n]).decode())
        sizes = [int(x) for x in head["b"]]
    except (ValueError, KeyError, TypeError) as e:
        raise WireFormatError(f"malformed header: {e!r}") from None
    off, blobs = 4 + n, []
    for sz in sizes:
        if sz < 0 or off + sz > len(buf):
            raise WireFormatError("blob runs past the frame")
        blobs.append(bytes(buf[off:off + sz]))
        off += sz
    return _unpack(head["t"], blobs)
