"""Device-visible micro-batch work queue: the B200 form of `DistributedDataset.next()` / `completeBatch()`
(`/root/reference/src/server/dataset.ts:43-67`, README "first come first serve" `/root/reference/README.md:33-45`).

The queue lives in the signal pad of a home rank's symmetric heap; any rank takes a ticket with a system-scope
atomic from inside a kernel (`wq_take_kernel`), gathers the batch rows of a device-resident dataset by that ticket
(`gather_batch_kernel`) and marks completion in a bitmap (`wq_complete_kernel`).  Properties kept:
FCFS index order, a batch counts only when completed, incomplete batches are re-issued when fresh ones run out,
the epoch advances only when all are complete, completion is idempotent.  The host never has to see a ticket;
`take()` (blocking read-back) exists for tests and logging, `take_async()` keeps it on the device.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from ..ops import _lib
from ..ops._lib import c_void_p, check, ptr
from .engine import _comm_sigs
from .symm import SymmetricHeap

__all__ = ["DeviceWorkQueue", "Ticket"]

_DONE = 0xFFFFFFFF


@dataclass
class Ticket:
    batch: int
    epoch: int
    done: bool
    dev: Optional[torch.Tensor] = None


class DeviceWorkQueue:
    def __init__(self, heap: SymmetricHeap, batches: int, epochs: int, home_rank: int = 0, init: bool = True):
        self.L = _comm_sigs()
        self.heap, self.batches, self.epochs, self.home = heap, batches, epochs, home_rank
        if batches <= self.L.dfb_wq_pad_capacity():
            off = self.L.dfb_wq_pad_offset_words() * 4                 # small datasets: the slot inside the signal pad
        else:
            # the completion bitmap is sized from the dataset (60 000 examples at batch 1 = 7.5 KB): a block of the heap,
            # allocated by every rank so that heap offsets stay identical across ranks
            off = heap.alloc(self.L.dfb_wq_words(batches) * 4)
        self.home_pad = c_void_p(heap.peer_bases[home_rank] + off)
        self.ticket = torch.zeros(2, dtype=torch.int32, device=heap.device)
        if init and heap.rank == home_rank:
            check(self.L.dfb_wq_init(self.home_pad, batches, epochs, _lib.stream()), "dfb_wq_init")
            torch.cuda.synchronize()

    @classmethod
    def local(cls, batches: int, epochs: int, device=None) -> "DeviceWorkQueue":
        return cls(SymmetricHeap(4096 + 4 * (32 + batches // 32), 0, 1, None, device), batches, epochs)

    def take_async(self) -> torch.Tensor:
        """Take a ticket into device memory (`[batch, epoch]`, batch = 0xffffffff when the dataset is exhausted)."""
        check(self.L.dfb_wq_take(self.home_pad, ptr(self.ticket), _lib.stream()), "dfb_wq_take")
        return self.ticket

    def take(self) -> Ticket:
        t = self.take_async().clone()
        b, e = (int(v) & 0xFFFFFFFF for v in t.tolist())
        return Ticket(b if b != _DONE else -1, e, b == _DONE, t)

    def complete(self, ticket: Optional[Ticket] = None, accepted: Optional[torch.Tensor] = None) -> None:
        """Mark the batch complete.  `accepted` (device uint32) makes it conditional: a batch whose gradient was refused by
        the staleness guard stays incomplete and is re-issued, like a dropped upload on the host server."""
        t = self.ticket if ticket is None or ticket.dev is None else ticket.dev
        check(self.L.dfb_wq_complete(self.home_pad, ptr(t), ptr(accepted), _lib.stream()), "dfb_wq_complete")

    def gather(self, data: torch.Tensor, batch_size: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Rows of the batch named by the last ticket, straight from the device-resident `data` (no host index)."""
        data2 = data.reshape(data.shape[0], -1)
        row_bytes = data2.shape[1] * data2.element_size()
        if out is None:
            out = torch.empty((batch_size, *data.shape[1:]), dtype=data.dtype, device=data.device)
        check(self.L.dfb_gather_batch(ptr(data2), ptr(out), ptr(self.ticket), row_bytes, batch_size, data2.shape[0],
                                      _lib.stream()), "dfb_gather_batch")
        return out
