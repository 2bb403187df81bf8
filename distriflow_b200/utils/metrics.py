"""Device-side metric ring + host reporters.

Replaces the reference's blocking `model.evaluate(...).dataSync()` before every step
(`/root/reference/src/client/asynchronousSGD_client.ts:50-53`, `/root/reference/src/common/models.ts:106-115`)
and its `console.log` logging (`/root/reference/src/server/abstract_server.ts:92-103`): the metrics kernel writes
`[matches, rows, loss]` into a slot of a device ring, the host copies finished slots asynchronously (pinned memory,
one step late) and can append them to a JSONL log; `ThroughputMeter` reports samples/s and gradient-bus GB/s.
"""
from __future__ import annotations

import ctypes
import json
import time
from typing import Any, Dict, List, Optional

import torch


class DeviceMetricRing:
    def __init__(self, slots: int = 64, device: Optional[torch.device] = None):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.slots = slots
        self.ring = torch.zeros(slots, 4, dtype=torch.float32, device=self.device)
        self.head = 0
        self._pending: List[Any] = []
        self.history: List[Dict[str, float]] = []

    def record(self, pred: torch.Tensor, labels: torch.Tensor, loss: Optional[torch.Tensor] = None) -> None:
        """Arg-max accuracy of `pred [rows, C]` against `labels` (indices or dense rows) + optional loss, on the device."""
        from ..ops import _lib
        from ..ops._lib import c_int, c_void_p, check, dt, ptr, stream
        L = _lib.lib()
        if not getattr(L, "_dfb_acc_sig", False):
            L.dfb_accuracy.restype = c_int
            L.dfb_accuracy.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, ctypes.c_long, ctypes.c_long, c_int,
                                       c_void_p]
            L._dfb_acc_sig = True
        slot = self.ring[self.head % self.slots]
        slot.zero_()
        p2 = pred.reshape(-1, pred.shape[-1]).contiguous()
        if labels.dtype in (torch.int64, torch.int32) and labels.dim() == pred.dim() - 1:
            lab, kind, ldl = labels.reshape(-1).contiguous(), (0 if labels.dtype == torch.int64 else 1), 0
        else:
            lab, kind, ldl = labels.reshape(-1, pred.shape[-1]).to(p2.dtype).contiguous(), 2, pred.shape[-1]
        check(L.dfb_accuracy(ptr(p2), ptr(lab), kind, ptr(slot), p2.shape[0], p2.shape[1], p2.stride(0), ldl, dt(p2), stream()),
              "dfb_accuracy")
        if loss is not None:
            slot[2:3].copy_(loss.detach().reshape(1).float())
        host = torch.empty(4, dtype=torch.float32, pin_memory=True)
        host.copy_(slot, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending.append((ev, host, self.head))
        self.head += 1

    def poll(self, block: bool = False) -> List[Dict[str, float]]:
        """Slots whose copies have completed (never blocks the device unless `block`)."""
        out = []
        while self._pending and (block or self._pending[0][0].query()):
            ev, host, idx = self._pending.pop(0)
            ev.synchronize()
            rows = max(float(host[1]), 1.0)
            out.append({"step": idx, "accuracy": float(host[0]) / rows, "rows": float(host[1]), "loss": float(host[2])})
        self.history.extend(out)
        return out


class ThroughputMeter:
    """samples/s and gradient-sync bus GB/s from CUDA events (device time)."""

    def __init__(self, samples_per_step: int, grad_bytes: int = 0, world: int = 1):
        self.samples, self.grad_bytes, self.world = samples_per_step, grad_bytes, world
        self.cuda = torch.cuda.is_available()            # host clock on a CPU-only box (plumbing configs)
        if self.cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
        self.steps, self._t0 = 0, 0.0

    def start(self) -> None:
        if self.cuda:
            self.e0.record()
        self._t0 = time.perf_counter()
        self.steps = 0

    def step(self) -> None:
        self.steps += 1

    def stop(self) -> Dict[str, float]:
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            ms = self.e0.elapsed_time(self.e1)
        else:
            ms = (time.perf_counter() - self._t0) * 1e3
        per = ms / max(self.steps, 1)
        bus = 2.0 * (self.world - 1) / self.world * self.grad_bytes / (per * 1e-3) / 1e9 if self.world > 1 else 0.0
        return {"ms_per_step": per, "samples_per_sec": self.samples * self.world / (per * 1e-3), "grad_bus_gbs": bus}


class JsonlLogger:
    def __init__(self, path: str):
        self.f = open(path, "a")

    def log(self, **kw: Any) -> None:
        kw.setdefault("ts", time.time())
        self.f.write(json.dumps(kw) + "\n")
        self.f.flush()

    def close(self) -> None:
        self.f.close()

    def __enter__(self) -> "JsonlLogger":
        return self

    def __exit__(self, *exc: Any) -> None:
        self.close()
