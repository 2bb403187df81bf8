"""DistriDataset: first-come-first-serve micro-batch dispatcher (host reference semantics).

Parity: `DistributedDataset` + `batchToDataMSG`
(`/root/reference/src/server/dataset.ts:5-109`) and the README design
(`/root/reference/README.md:33-45`: epoch IDs, micro-batch IDs, queued /
incomplete micro-batches, queued / completed epochs).

Properties kept:
  1. batches are issued FCFS in index order;
  2. a batch counts for the epoch only when its gradient came back
     (`complete_batch`);
  3. batches issued but not completed are re-issued once the fresh ones run
     out (straggler / lost-worker recovery);
  4. the epoch advances only when every batch of the epoch is complete;
  5. duplicates are possible and `complete_batch` is idempotent.

Fixed relative to the reference: `smallLastBatch=True` yields a
short final batch, `False` drops the remainder (the reference stores the flag
and then throws in `slice` on a ragged tail, dataset.ts:27,76-77).  A
completion that names a previous epoch is ignored instead of deleting the
same index from the new epoch's set.

The device-side counterpart (atomic ticket + completion bitmap in symmetric
memory) is `parallel/workqueue.py`; it is tested against this class.
"""
from __future__ import annotations

import math
import uuid
from typing import Any, Callable, Dict, List, Optional

import torch

from .config import dataset_config
from .messages import Batch, DataMsg
from .serialization import serialize_var

__all__ = ["DistriDataset", "DistributedDataset", "batch_to_data_msg"]


class DistriDataset:
    def __init__(self, x: torch.Tensor, y: torch.Tensor, config: Optional[Dict[str, Any]] = None,
                 *, use_uuids: bool = True):
        if x.shape[0] != y.shape[0]:
            raise ValueError(f"x has {x.shape[0]} rows, y has {y.shape[0]}")
        cfg = dataset_config(config)
        self.x, self.y = x, y
        self.epochs = int(cfg["epochs"])
        self.batchSize = int(cfg["batchSize"])
        self.smallLastBatch = bool(cfg["smallLastBatch"])
        n = int(x.shape[0])
        self.batches = math.ceil(n / self.batchSize) if self.smallLastBatch else n // self.batchSize
        if self.batches == 0:
            raise ValueError("dataset smaller than one batch and smallLastBatch is False")
        self.epoch = 0
        self.preprocessCallbacks: List[Callable[[Batch], Batch]] = []
        self._use_uuids = use_uuids
        # README.md:33-45 bookkeeping
        self.epochIDs: List[str] = [self._new_id() for _ in range(self.epochs)]
        self.CompletedEpochs: List[str] = []
        self.issued: Dict[int, int] = {}          # batch -> times issued in this epoch
        self.duplicates = 0                        # completions that arrived twice
        self._start_epoch()

    # ---- ids -------------------------------------------------------------------------------
    def _new_id(self) -> str:
        return str(uuid.uuid4()) if self._use_uuids else ""

    @property
    def QueuedEpochs(self) -> List[str]:
        return self.epochIDs[self.epoch + 1:]

    @property
    def epochID(self) -> Optional[str]:
        return self.epochIDs[self.epoch] if self.epoch < self.epochs else None

    @property
    def incompleteBatches(self) -> List[int]:
        """Indices still without a gradient, in issue order (dataset.ts:31)."""
        return [b for b in self._order if b in self._incomplete]

    incompletedMicroBatches = incompleteBatches

    @property
    def queuedMicroBatch(self) -> List[int]:
        """Not-yet-issued batches of the current pass."""
        return [b for b in self._order[self._cursor:] if b in self._incomplete]

    # ---- dispatcher -------------------------------------------------------------------------
    def _start_epoch(self) -> None:
        self._order = list(range(self.batches))
        self._incomplete = set(self._order)
        self._cursor = 0
        self.microBatchIDs = [self._new_id() for _ in range(self.batches)]
        self.issued = {}

    def complete_batch(self, batch: Optional[int], epoch: Optional[int] = None) -> bool:
        """Mark a batch done; idempotent (dataset.ts:43-45)."""
        if batch is None:
            return False
        if epoch is not None and epoch != self.epoch:
            return False
        if batch in self._incomplete:
            self._incomplete.discard(batch)
            return True
        self.duplicates += 1
        return False

    completeBatch = complete_batch

    def _pull(self) -> Optional[int]:
        while self._cursor < len(self._order):
            b = self._order[self._cursor]
            self._cursor += 1
            if b in self._incomplete:     # live iterator: entries deleted meanwhile are skipped
                return b
        return None

    def next(self) -> Batch:
        """Next batch to hand out, or `Batch(done=True)` after the last epoch (dataset.ts:47-67)."""
        if self.epoch >= self.epochs:
            return Batch(-1, self.epoch, None, None, done=True)
        if not self._incomplete:
            self.CompletedEpochs.append(self.epochIDs[self.epoch])
            self.epoch += 1
            if self.epoch >= self.epochs:
                return Batch(-1, self.epoch, None, None, done=True)
            self._start_epoch()
        b = self._pull()
        if b is None:                     # fresh ones ran out: re-issue what is still incomplete
            self._order = [i for i in self._order if i in self._incomplete]
            self._cursor = 0
            b = self._pull()
        assert b is not None
        self.issued[b] = self.issued.get(b, 0) + 1
        return self._preprocess(self.get_batch(b))

    def get_batch(self, batch: int) -> Batch:
        lo = batch * self.batchSize
        hi = min(lo + self.batchSize, int(self.x.shape[0]))
        return Batch(batch, self.epoch, self.x[lo:hi], self.y[lo:hi],
                     batchId=self.microBatchIDs[batch], epochId=self.epochID)

    def _preprocess(self, batch: Batch) -> Batch:
        for cb in self.preprocessCallbacks:
            batch = cb(batch)
        return batch

    def add_preprocess_callback(self, cb: Callable[[Batch], Batch]) -> None:
        self.preprocessCallbacks.append(cb)

    addPreprocessCallback = add_preprocess_callback

    # ---- iteration helpers -----------------------------------------------------------------
    def __iter__(self):
        while True:
            b = self.next()
            if b.done:
                return
            yield b

    @property
    def done(self) -> bool:
        return self.epoch >= self.epochs or (self.epoch == self.epochs - 1 and not self._incomplete)

    def state_dict(self) -> Dict[str, Any]:
        return {"epoch": self.epoch, "incomplete": sorted(self._incomplete), "cursor": self._cursor,
                "order": list(self._order)}

    def load_state_dict(self, st: Dict[str, Any]) -> None:
        self.epoch = int(st["epoch"])
        self._order = list(st["order"])
        self._incomplete = set(st["incomplete"])
        self._cursor = int(st["cursor"])


DistributedDataset = DistriDataset   # reference class name


def batch_to_data_msg(batch: Batch) -> DataMsg:
    """Serialise a batch for a host transport (dataset.ts:99-109)."""
    return DataMsg(batch=batch.batch, epoch=batch.epoch, x=serialize_var(batch.x), y=serialize_var(batch.y),
                   batchId=batch.batchId, epochId=batch.epochId)


batchToDataMSG = batch_to_data_msg
