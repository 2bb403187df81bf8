"""Training loop around the data-parallel engine: epochs, micro-batches, versioned checkpoints, resume.

This is the GPU-rank counterpart of what the reference's server does around its model (`/root/reference/src/server/
asynchronousSGD_server.ts:65-108`: dispatch a micro-batch, apply the gradient, `model.save()`, broadcast) with the
reference's checkpoint layout (`/root/reference/src/server/models.ts:98-150`: `<saveDir>/<version>/model.json` +
`weights.bin`, a `current` link, resume from the newest version on set-up).  Differences, all deliberate:

* every rank holds the dataset (or its shard) and takes ITS slice of each global batch — a synchronous step needs no
  dispatcher; the first-come-first-serve dispatcher with re-issue lives on in `DistriDataset` (host servers) and in the
  device work queue (async engine);
* a version directory is written every `checkpointEvery` steps, not after every update, and holds
  `engine_state.pt` (optimizer moments, step count) next to the tf.js artifact, so a resumed run continues bit for bit;
* the learning rate may follow a schedule (`lrSchedule(step) -> lr`), which also reaches steps replayed from a CUDA graph.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from .checkpoint import force_symlink, list_versions, new_version
from .config import ConfigError, override

__all__ = ["Trainer", "DEFAULT_TRAINER_CONFIG"]

DEFAULT_TRAINER_CONFIG: Dict[str, Any] = {
    "batchSize": 32,              # per rank (the reference's micro-batch size, utils.ts:193-197)
    "epochs": 5,
    "smallLastBatch": False,      # False: drop the ragged tail (shapes stay fixed: required under cudaGraph)
    "saveDir": None,              # version directories + `current` link; None = no checkpoints
    "checkpointEvery": 0,         # steps between checkpoints (0 = only at the end when saveDir is set)
    "keepVersions": 0,            # > 0: delete all but the newest N version directories
    "resume": True,               # continue from `<saveDir>/current` when it exists
    "logPath": None,              # JSONL log of (step, epoch, batch, loss, lr)
    "logEvery": 1,
}


class Trainer:
    """`Trainer(engine, x, y, config).fit()`.

    `x`, `y`: tensors holding this rank's view of the dataset; with `shard=True` (default) every rank holds the WHOLE
    dataset and uses rows `[g*W*B + r*B, g*W*B + (r+1)*B)` of global batch `g` (W ranks, batch size B); with
    `shard=False` the tensors are already this rank's private shard.
    """

    def __init__(self, engine: Any, x: torch.Tensor, y: torch.Tensor, config: Optional[Dict[str, Any]] = None, *,
                 shard: bool = True, lr_schedule: Optional[Callable[[int], float]] = None):
        try:
            self.cfg = override(DEFAULT_TRAINER_CONFIG, config)
        except ConfigError as err:
            raise ConfigError(f"Error setting trainerConfig: {err}") from None
        self.engine, self.x, self.y, self.lr_schedule = engine, x, y, lr_schedule
        self.rank, self.world = int(getattr(engine, "rank", 0)), int(getattr(engine, "world", 1))
        self.bs = int(self.cfg["batchSize"])
        if self.bs < 1 or int(self.cfg["epochs"]) < 0:
            raise ConfigError("Error setting trainerConfig: batchSize >= 1 and epochs >= 0 required")
        self.shard = bool(shard) and self.world > 1
        n = int(x.shape[0])
        per_step = self.bs * (self.world if self.shard else 1)
        self.batches = n // per_step if not self.cfg["smallLastBatch"] else -(-n // per_step)
        if self.batches < 1:
            raise ValueError(f"dataset of {n} examples is smaller than one global batch of {per_step}")
        self.per_step = per_step
        self.stepCallbacks: List[Callable[[int, int, int, torch.Tensor], None]] = []
        self.checkpointCallbacks: List[Callable[[str], None]] = []
        self.saved: List[str] = []
        self._log = None
        if self.cfg["logPath"] and self.rank == 0:
            from .utils.metrics import JsonlLogger
            self._log = JsonlLogger(self.cfg["logPath"])

    # ------------------------------------------------------------------------------------------ callbacks
    def on_step(self, cb: Callable[[int, int, int, torch.Tensor], None]) -> None:
        """`cb(step, epoch, batch, loss)` after every step (loss is a device tensor: reading it synchronises)."""
        self.stepCallbacks.append(cb)

    def on_checkpoint(self, cb: Callable[[str], None]) -> None:
        self.checkpointCallbacks.append(cb)

    # ------------------------------------------------------------------------------------------ data
    def batch(self, index: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """This rank's slice of global batch `index` of an epoch."""
        lo = index * self.per_step + (self.rank * self.bs if self.shard else 0)
        hi = min(lo + self.bs, int(self.x.shape[0]))
        return self.x[lo:hi], self.y[lo:hi]

    # ------------------------------------------------------------------------------------------ checkpoints
    def save(self) -> Optional[str]:
        """New version directory: tf.js artifact of the fp32 masters + engine state; `current` is moved last.  Collective."""
        save_dir = self.cfg["saveDir"]
        if not save_dir:
            return None
        # A step that averaged incomplete gradients (a peer GPU was lost: `PeerLostError`) must never become a resume point:
        # the last good version stays `current`, a supervisor restarts the job on the surviving ranks and `fit()` resumes
        # from it (the state is independent of the world size).  Reference: a worker may leave at any time and the lost
        # batches are re-issued (/root/reference/src/server/dataset.ts:57-60, asynchronousSGD_server.ts:54-57).
        if hasattr(self.engine, "check_health"):
            self.engine.check_health()
        version = None
        if self.rank == 0:
            os.makedirs(save_dir, exist_ok=True)
            version = new_version()
        if self.world > 1:
            import torch.distributed as dist
            box = [version]
            dist.broadcast_object_list(box, src=0, group=getattr(self.engine, "group", None))
            version = box[0]
        path = os.path.join(save_dir, version)
        self.engine.save_checkpoint(path)
        self.engine.save_state(path)
        if self.rank != 0:
            return None
        force_symlink(version, os.path.join(save_dir, "current"))          # relative link, like the reference
        self.saved.append(path)
        keep = int(self.cfg["keepVersions"])
        if keep > 0:
            import shutil
            for old in list_versions(save_dir)[:-keep]:
                shutil.rmtree(os.path.join(save_dir, old), ignore_errors=True)
        for cb in self.checkpointCallbacks:
            cb(path)
        return path

    def latest(self) -> Optional[str]:
        save_dir = self.cfg["saveDir"]
        if not save_dir:
            return None
        cur = os.path.join(save_dir, "current")
        if os.path.exists(os.path.join(cur, "engine_state.pt")):
            return cur
        for v in reversed(list_versions(save_dir)):                       # a crash between the directory and the link
            if os.path.exists(os.path.join(save_dir, v, "engine_state.pt")):
                return os.path.join(save_dir, v)
        return None

    # ------------------------------------------------------------------------------------------ loop
    def fit(self, max_steps: Optional[int] = None) -> Dict[str, Any]:
        """Run to the configured number of epochs (or `max_steps` more steps).  Returns a summary dict."""
        eng = self.engine
        if self.cfg["resume"]:
            path = self.latest()
            if path is not None:
                eng.load_state(path)
        total = self.batches * int(self.cfg["epochs"])
        step = int(eng.version)                              # steps already applied = position in the data stream
        start, done_now, last_loss = step, 0, None
        every = int(self.cfg["checkpointEvery"])
        dev = getattr(eng, "device", torch.device("cpu"))

        def stream_of_batches():
            s = step
            while s < total:
                yield self.batch(s % self.batches)
                s += 1

        from .utils.prefetch import DevicePrefetcher
        feed = DevicePrefetcher(stream_of_batches(), dev, depth=2)      # pass-through on CPU; H2D on a copy stream on CUDA
        while step < total and (max_steps is None or done_now < max_steps):
            epoch, b = divmod(step, self.batches)
            if self.lr_schedule is not None:
                eng.set_learning_rate(float(self.lr_schedule(step)))
            xb, yb = next(feed)
            last_loss = eng.step(xb, yb)
            step += 1
            done_now += 1
            for cb in self.stepCallbacks:
                cb(step, epoch, b, last_loss)
            if self._log is not None and step % max(int(self.cfg["logEvery"]), 1) == 0:
                self._log.log(step=step, epoch=epoch, batch=b, loss=float(last_loss), lr=float(eng.lr))
            if every > 0 and step % every == 0:
                self.save()
        finished = step >= total
        just_saved = every > 0 and done_now > 0 and step % every == 0
        if self.cfg["saveDir"] and done_now > 0 and not just_saved:      # the stopping point is always a resume point
            self.save()
        if self._log is not None and finished:
            self._log.close()
            self._log = None
        return {"steps": step, "steps_this_call": done_now, "resumed_from": start, "finished": finished,
                "loss": None if last_loss is None else float(last_loss), "checkpoints": list(self.saved)}
