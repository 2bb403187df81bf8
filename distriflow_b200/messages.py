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

import pickle
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional

from .serialization import SerializedVariable

__all__ = ["Events", "ModelMsg", "GradientMsg", "DataMsg", "Batch", "UploadMsg",
           "DownloadMsg", "VersionCallback", "UploadCallback", "PreprocessCallback",
           "encode", "decode"]


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


def encode(obj: Any) -> bytes:
    """Binary framing payload for the TCP transport."""
    return pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)


def decode(buf: bytes) -> Any:
    return pickle.loads(buf)
