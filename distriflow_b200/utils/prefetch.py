"""Host -> device input prefetch on a copy stream.

The reference ships every micro-batch through the websocket inside the `downloadVars` message and deserialises it on the
worker before the step (`/root/reference/src/server/dataset.ts:69-85`, `/root/reference/src/client/
asynchronousSGD_client.ts:46-47`): the transfer is on the critical path of every step.  Here the next batches are copied
from pinned host memory on a side stream while the current step computes; the consumer stream waits on the copy's event,
never on the host.  On a CPU device the prefetcher is a plain pass-through, so the same training loop runs everywhere.
"""
from __future__ import annotations

import collections
from typing import Any, Deque, Iterable, Iterator, Optional, Sequence, Tuple

import torch

__all__ = ["DevicePrefetcher"]


class DevicePrefetcher:
    """Iterate `(tensor, ...)` tuples of an iterable with `depth` batches already on their way to `device`.

    Host tensors are pinned on first sight if they are not already (pinning is what makes the copy asynchronous); tensors
    that already live on `device` pass through untouched.
    """

    def __init__(self, batches: Iterable[Sequence[Any]], device: torch.device, depth: int = 2, pin: bool = True):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.depth, self.pin = depth, pin
        self._it: Iterator[Sequence[Any]] = iter(batches)
        self._q: Deque[Tuple[Tuple[Any, ...], Optional[Any]]] = collections.deque()
        self._stream = torch.cuda.Stream(device=self.device) if self.cuda else None
        self.h2d_bytes = 0

    def _stage(self) -> bool:
        try:
            batch = next(self._it)
        except StopIteration:
            return False
        if not self.cuda:
            self._q.append((tuple(batch), None))
            return True
        out = []
        with torch.cuda.stream(self._stream):
            for t in batch:
                if isinstance(t, torch.Tensor) and t.device != self.device:
                    if self.pin and t.device.type == "cpu" and not t.is_pinned():
                        t = t.pin_memory()
                    self.h2d_bytes += t.numel() * t.element_size()
                    t = t.to(self.device, non_blocking=True)
                out.append(t)
            ev = torch.cuda.Event()
            ev.record(self._stream)
        self._q.append((tuple(out), ev))
        return True

    def __iter__(self) -> "DevicePrefetcher":
        return self

    def __next__(self) -> Tuple[Any, ...]:
        while len(self._q) < self.depth and self._stage():
            pass
        if not self._q:
            raise StopIteration
        batch, ev = self._q.popleft()
        if ev is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)                                   # device-side wait: the host does not block
            for t in batch:
                if isinstance(t, torch.Tensor):
                    t.record_stream(cur)                         # the allocator must not recycle it under the consumer
        self._stage()                                            # keep the pipeline full while the caller computes
        return batch
