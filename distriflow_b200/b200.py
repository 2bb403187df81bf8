"""`DistriServer(..., backend="b200")` / `DistriWorker(..., backend="b200")`: the reference's object model on the GPU path.

In the reference the server object IS the training loop: it owns the model and the dataset, hands out micro-batches,
applies uploads, broadcasts versions, counts clients and fires `onUpload` / `onNewVersion`
(`/root/reference/src/server/asynchronousSGD_server.ts:45-82`, `/root/reference/src/server/abstract_server.ts:67-115`,
README design `/root/reference/README.md:8-45`).  On one 8xB200 box the same four objects exist with the roles the
north star gives them: every worker is a GPU rank, the server is SHARDED across the ranks, so every rank process builds
the same `DistriServer` (SPMD) and drives its own share of it:

  DistriServer   model + dataset + `workerIDs` (one id per rank, exchanged at bring-up), `numClients`, `numUpdates`,
                 `numDropped`, `on_upload(msg)` / `on_new_version(old, new)`, `saveEvery` checkpoints in the tf.js layout.
                 mode "sync"  -> `parallel.DataParallelEngine` (fused reduce-scatter + optimizer + all-gather kernels);
                 `minUpdatesPerVersion = K x workers` accumulates K micro-batches per rank into the GPU-resident
                 `savedGradient` before ONE fused kernel (`/root/reference/src/server/federated_server.ts:73-90`).
                 mode "async" -> `parallel.AsyncEngine` (P2P push behind the device-side `maximumStaleness` guard).
  DistriDataset  its config (batchSize, epochs, smallLastBatch) sizes the device work queue; x / y live on the GPU and
                 batches are gathered by ticket (`/root/reference/src/server/dataset.ts:47-67` FCFS + re-issue).
  DistriModel    `server.model`: version (= optimizer steps), `maximumStaleness`, `savedGradient` (the engine's buckets).
  DistriWorker   `B200Worker`: this rank's handle (clientId, counters, `run()`), cf. `abstract_client.ts:80-122`.

Per-upload worker metrics (`sendMetrics`, `/root/reference/src/client/asynchronousSGD_client.ts:50-53,77-79`): loss and
arg-max accuracy of the batch on the weights the gradient was computed with are written by `accuracy_kernel` into the
`DeviceMetricRing`; `on_upload` receives them in `UploadMsg.metrics` a few steps late, never blocking the GPU.

On a machine without CUDA the same classes run the engine's gloo/CPU back end with the host dispatcher, which is how the
reference's three API scenarios are tested here (`tests/test_b200_api.py`).
"""
from __future__ import annotations

import math
import os
import uuid
from typing import Any, Callable, Dict, List, Optional, Set

import torch

from .config import client_hyperparams, compile_args, server_hyperparams
from .dataset import DistriDataset
from .losses import get_loss
from .messages import UploadMsg

__all__ = ["B200Server", "B200Worker", "B200Model"]

_CONFIG_KEYS = {"clientHyperparams", "serverHyperparams", "modelDir", "modelCompileConfig", "verbose", "saveEvery",
                "sendMetrics", "engine", "clientId"}


class B200Model:
    """`DistriModel` view of the engine: version / modelID, maximumStaleness, savedGradient (README.md:23-30)."""

    isDistributedModel = True

    def __init__(self, server: "B200Server"):
        self._s = server

    @property
    def module(self) -> torch.nn.Module:
        return self._s.module

    @property
    def versionIndex(self) -> int:
        return int(self._s.engine.version) if self._s.engine is not None else 0

    @property
    def version(self) -> str:
        return str(self.versionIndex)

    modelID = version

    @property
    def maximumStaleness(self) -> int:
        return int(self._s.serverHyperparams["maximumStaleness"])

    @property
    def savedGradient(self) -> Optional[List[torch.Tensor]]:
        return getattr(self._s.engine, "savedGradient", None)

    @property
    def savedCount(self) -> int:
        return int(getattr(self._s.engine, "savedCount", 0))

    def get_vars(self) -> List[torch.Tensor]:
        eng = self._s.engine
        if hasattr(eng, "full_precision_weights"):
            return eng.full_precision_weights()
        return [p.detach().float() for p in self._s.module.parameters() if p.requires_grad]

    getVars = get_vars

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            return self._s.module(x.to(self._s.device))


class B200Server:
    def __init__(self, model: Any, dataset: Optional[DistriDataset] = None, config: Optional[Dict[str, Any]] = None,
                 mode: str = "sync", rank: Optional[int] = None, world: Optional[int] = None, group=None,
                 device: Optional[torch.device] = None):
        config = dict(config or {})
        for k in config:
            if k not in _CONFIG_KEYS:
                raise ValueError(f'Unrecognized server config key "{k}"')
        if mode not in ("sync", "async", "federated"):
            raise ValueError(f'backend "b200" runs modes "sync", "async" and "federated", got {mode!r}')
        if dataset is None and mode != "federated":
            raise ValueError("the b200 back end owns the dataset (DistriServer {model, dataset, workerIDs}); only federated "
                             "workers bring their own data")
        self.mode, self.dataset, self.group = mode, dataset, group
        self.clientHyperparams = client_hyperparams(config.get("clientHyperparams"))
        shp = dict(config.get("serverHyperparams") or {})
        if mode == "async":
            shp.setdefault("maximumStaleness", 4)
        self.serverHyperparams = server_hyperparams(shp)
        self.compile = compile_args(config.get("modelCompileConfig"))
        self.engine_overrides = dict(config.get("engine") or {})
        self.modelDir = config.get("modelDir")
        self.saveEvery = max(1, int(config.get("saveEvery") or 1))
        self.sendMetrics = bool(config.get("sendMetrics"))
        verbose = config.get("verbose")
        self.verbose = bool(os.environ.get("VERBOSE")) if verbose is None else bool(verbose)
        import torch.distributed as dist
        ddp = dist.is_available() and dist.is_initialized()
        self.rank = int(rank if rank is not None else (dist.get_rank(group) if ddp else 0))
        self.world = int(world if world is not None else (dist.get_world_size(group) if ddp else 1))
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.device = device
        self.clientId = str(config.get("clientId") or f"gpu{self.rank}-{uuid.uuid4().hex[:8]}")
        self._model_arg = model
        self.module: Optional[torch.nn.Module] = None
        self.engine: Any = None
        self.model = B200Model(self)
        self.workerIDs: Set[str] = set()
        self.numClients = 0
        self.numUpdates = 0               # gradients this rank contributed (uploads)
        self.numDropped = 0               # async: gradients refused by the staleness guard (whole job, from the owners' counters)
        self.versionCallbacks: List[Callable[[Optional[str], str], None]] = []
        self.uploadCallbacks: List[Callable[[Optional[UploadMsg]], None]] = []
        self.done = False
        self.lastLoss: Optional[torch.Tensor] = None
        self.ring = None
        self.queue = None
        self._since_save = 0
        self._out = None

    # hooks (abstract_server.ts:67-79) ---------------------------------------------------------
    def on_new_version(self, cb: Callable[[Optional[str], str], None]) -> None:
        self.versionCallbacks.append(cb)

    def on_upload(self, cb: Callable[[Optional[UploadMsg]], None]) -> None:
        self.uploadCallbacks.append(cb)

    onNewVersion, onUpload = on_new_version, on_upload

    def log(self, *args: Any) -> None:
        if self.verbose:
            print(f"Distributed Server[{self.rank}]:", *args, flush=True)

    # bring-up ---------------------------------------------------------------------------------
    def setup(self) -> None:
        """Resolve the model, register the workers, build the engine (server shards) and the dispatcher."""
        from .checkpoint import fetch_model
        m = self._model_arg
        loss_name = self.compile["loss"]
        if getattr(m, "isDistributedModel", False) and hasattr(m, "_resolve"):          # a DistriTorchModel wrapper
            m._resolve()
            self.module, loss_name = m.model, m.loss
            lr = float(m.learningRate)
        else:
            self.module = fetch_model(m)
            lr = float(self.compile["learningRate"])
        self.module.to(self.device)
        lossf = get_loss(loss_name)

        def loss_fn(mod, x, y):
            out = mod(x)
            self._out = (out, y)                   # the batch metrics of this upload are taken from the training forward
            return lossf(y, out)

        self.workerIDs = {self.clientId}
        if self.world > 1:
            import torch.distributed as dist
            ids: List[Optional[str]] = [None] * self.world
            dist.all_gather_object(ids, self.clientId, group=self.group)
            self.workerIDs = set(ids)
        self.numClients = len(self.workerIDs)
        self.microBatchesPerVersion = 1
        cuda = self.device.type == "cuda"
        ecfg = {"optimizer": "sgd", "learningRate": lr, "maximumStaleness": int(self.serverHyperparams["maximumStaleness"])}
        if not cuda or next(self.module.parameters()).dtype == torch.float32:
            ecfg.update({"dtype": "fp32", "gradDtype": "fp32"})
        ecfg.update(self.engine_overrides)
        if self.mode in ("sync", "federated"):
            # federated (federated_server.ts:71-116): `minUpdatesPerVersion` uploads are buffered and averaged into one version;
            # here every rank contributes the same number of them, from its OWN examples (`B200Worker.federated_update`)
            from .parallel.engine import DataParallelEngine
            self.microBatchesPerVersion = max(1, math.ceil(int(self.serverHyperparams["minUpdatesPerVersion"]) / self.world))
            if self.microBatchesPerVersion > 1:
                ecfg["cudaGraph"] = False
            self.engine = DataParallelEngine(self.module, loss_fn, {**ecfg, "mode": "sync"}, rank=self.rank, world=self.world,
                                             group=self.group, device=self.device)
        else:
            if not cuda:
                raise RuntimeError('mode "async" of the b200 back end needs CUDA devices (peer-memory pushes); use the host servers on CPU')
            from .parallel.async_engine import AsyncEngine
            self.engine = AsyncEngine(self.module, loss_fn, {**ecfg, "mode": "async"}, rank=self.rank, world=self.world,
                                      group=self.group, device=self.device)
        ds = self.dataset
        if ds is None:                                  # federated: no server-owned data, no dispatcher
            self.batchSize, self.batches, self.epochs = int(self.clientHyperparams["examplesPerUpdate"]), 0, 0
            if cuda and self.sendMetrics:
                from .utils.metrics import DeviceMetricRing
                self.ring = DeviceMetricRing(slots=64, device=self.device)
            self._lag = []
            self._last_host_batch = None
            self.log(f"setup: {self.numClients} federated workers, {self.microBatchesPerVersion} uploads per rank and version")
            return
        self.batchSize, self.batches, self.epochs = ds.batchSize, ds.batches, ds.epochs
        self.x_all, self.y_all = ds.x.to(self.device), ds.y.to(self.device)
        if cuda:
            from .parallel.symm import SymmetricHeap
            from .parallel.workqueue import DeviceWorkQueue
            from .parallel.engine import _comm_sigs
            L = _comm_sigs()
            heap = SymmetricHeap(4096 + 4 * int(L.dfb_wq_words(self.batches)), self.rank, self.world, self.group, self.device)
            self.queue = DeviceWorkQueue(heap, self.batches, self.epochs, home_rank=0)
            self._xb = torch.empty((self.batchSize, *self.x_all.shape[1:]), dtype=self.x_all.dtype, device=self.device)
            self._yb = torch.empty((self.batchSize, *self.y_all.shape[1:]), dtype=self.y_all.dtype, device=self.device)
            if self.sendMetrics:
                from .utils.metrics import DeviceMetricRing
                self.ring = DeviceMetricRing(slots=64, device=self.device)
            if self.world > 1:
                import torch.distributed as dist
                torch.cuda.synchronize()
                dist.barrier(group=self.group)      # the queue is initialised on its home rank before anybody takes a ticket
        self._lag: List[Any] = []
        self._last_host_batch = None
        self.log(f"setup: {self.numClients} workers, {self.batches} micro-batches x {self.epochs} epochs, mode {self.mode}, "
                 f"{self.microBatchesPerVersion} micro-batches per rank and version")

    # one upload -------------------------------------------------------------------------------
    def _next_batch(self):
        """(x, y, ticket) of the next micro-batch of this rank; ticket is a device tensor on CUDA, a host Batch on CPU."""
        if self.queue is not None:
            t = self.queue.take_async()
            self.queue.gather(self.x_all, self.batchSize, self._xb)
            self.queue.gather(self.y_all, self.batchSize, self._yb)
            return self._xb, self._yb, t
        mine = None
        for r in range(self.world):               # every replica of the host dispatcher sees the same sequence of events
            b = self.dataset.next()
            if b.done:
                self.done = True
                break
            self.dataset.complete_batch(b.batch, b.epoch)
            if r == self.rank:
                mine = b
        if mine is None:
            mine = self._last_host_batch
        self._last_host_batch = mine
        return mine.x.to(self.device), mine.y.to(self.device), mine

    def _uploaded(self, ticket: Any, loss: torch.Tensor) -> None:
        """Bookkeeping of one contributed gradient: counters, metrics into the device ring, `on_upload` (lagging, non-blocking)."""
        self.numUpdates += 1
        self.lastLoss = loss
        if self.queue is None:                                          # CPU plumbing: everything is already on the host
            metrics = None
            if self.sendMetrics and self._out is not None:
                out, y = self._out
                lab = y if y.dim() == out.dim() - 1 else y.argmax(-1)
                metrics = [float(loss), float((out.detach().argmax(-1) == lab).float().mean())]
            self._fire_upload(ticket.batch, ticket.epoch, metrics)
            return
        if not self.uploadCallbacks:
            return
        if self.ring is not None and self._out is not None:
            out, y = self._out
            self.ring.record(out.detach(), y, loss)
        host = torch.empty(2, dtype=torch.int32).pin_memory()
        host.copy_(ticket, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._lag.append((ev, host))
        self._poll(block=False)

    def _poll(self, block: bool) -> None:
        ready = self.ring.poll(block=block) if self.ring is not None else None
        if ready is not None:
            self._ring_ready = getattr(self, "_ring_ready", []) + ready
        while self._lag and (block or self._lag[0][0].query()):
            ev, host = self._lag.pop(0)
            ev.synchronize()
            b, e = int(host[0]) & 0xFFFFFFFF, int(host[1])
            metrics = None
            if self.ring is not None:
                if not getattr(self, "_ring_ready", None):
                    self._ring_ready = self.ring.poll(block=True)
                rec = self._ring_ready.pop(0)
                metrics = [rec["loss"], rec["accuracy"]]
            if b != 0xFFFFFFFF:
                self._fire_upload(b, e, metrics)

    def _fire_upload(self, batch: Optional[int], epoch: Optional[int], metrics: Optional[List[float]]) -> None:
        msg = UploadMsg(clientId=self.clientId, gradients=None, batch=batch, metrics=metrics, epoch=epoch)
        for cb in self.uploadCallbacks:
            cb(msg)

    def _new_version(self, old: int) -> None:
        for cb in self.versionCallbacks:
            cb(str(old), self.model.version)
        self._since_save += 1
        if self.modelDir and self._since_save >= self.saveEvery and hasattr(self.engine, "save_checkpoint"):
            self._since_save = 0
            from .checkpoint import force_symlink, new_version
            name = new_version()
            path = self.engine.save_checkpoint(os.path.join(self.modelDir, name))       # collective; rank 0 writes
            if path is not None:
                force_symlink(os.path.join(self.modelDir, name), os.path.join(self.modelDir, "current"))

    # the training loop ------------------------------------------------------------------------
    @property
    def total_uploads_per_rank(self) -> int:
        """Micro-batches every rank processes in a synchronous run: the dataset's tickets shared out evenly (the collective
        kernels need every rank in every version; a rank whose last ticket says "exhausted" repeats its previous batch)."""
        return math.ceil(self.batches * self.epochs / self.world)

    def upload(self) -> Optional[torch.Tensor]:
        """Synchronous mode, ONE micro-batch of this rank (every rank calls it the same number of times): take a ticket,
        gather the batch, compute the gradient.  The first K-1 gradients of a version are added to the GPU-resident
        `savedGradient`; the K-th closes the version: mean over the K micro-batches and all ranks, applied by the fused
        reduce-scatter + optimizer + all-gather kernels, `on_new_version` fires.  Returns the batch loss, None when done."""
        if self.mode != "sync":
            raise RuntimeError("upload()/step() drive the synchronous server; the asynchronous one runs with run()")
        if self.engine is None:
            self.setup()
        if self.numUpdates >= self.total_uploads_per_rank:
            self.done = True
            return None
        x, y, ticket = self._next_batch()
        last_of_run = self.numUpdates + 1 >= self.total_uploads_per_rank
        closes = self.engine.savedCount + 1 >= self.microBatchesPerVersion or last_of_run    # a partial buffer is flushed at the end
        old = int(self.engine.version)
        loss = self.engine.step(x, y) if closes else self.engine.accumulate(x, y)
        if self.queue is not None:
            self.queue.complete()
        self._uploaded(ticket, loss)
        if closes:
            self._new_version(old)
        return loss

    def federated_upload(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Federated mode, ONE upload of this rank from the worker's own examples (collective when world > 1: every rank
        uploads the same number of times per version).  `minUpdatesPerVersion` uploads over all ranks make a version:
        the earlier ones are summed into `savedGradient` on the GPU, the closing one runs the fused mean + update kernels
        (/root/reference/src/server/federated_server.ts:71-116; client side federated_client.ts:85-128)."""
        if self.mode != "federated":
            raise RuntimeError('federated_upload() needs mode "federated"')
        if self.engine is None:
            self.setup()
        x, y = x.to(self.device), y.to(self.device)
        closes = self.engine.savedCount + 1 >= self.microBatchesPerVersion
        old = int(self.engine.version)
        loss = self.engine.step(x, y) if closes else self.engine.accumulate(x, y)
        self.numUpdates += 1
        self.lastLoss = loss
        metrics = None
        if self.sendMetrics and self._out is not None:
            out, yy = self._out
            lab = yy if yy.dim() == out.dim() - 1 else yy.argmax(-1)
            metrics = [float(loss), float((out.detach().argmax(-1) == lab).float().mean())]
        if self.uploadCallbacks:
            self._fire_upload(None, None, metrics)
        if closes:
            self._new_version(old)
        return loss

    def step(self) -> Optional[torch.Tensor]:
        """Uploads until the next model version exists; returns the loss of its last micro-batch (None when done)."""
        if self.engine is None:
            self.setup()
        start, loss = int(self.engine.version), None
        while int(self.engine.version) == start:
            loss = self.upload()
            if loss is None:
                return None
        return loss

    def run(self, max_versions: Optional[int] = None) -> Dict[str, Any]:
        """Train until the dataset's epochs are exhausted (or `max_versions` new versions were produced)."""
        if self.engine is None:
            self.setup()
        start = int(self.engine.version)
        if self.mode == "federated":
            raise RuntimeError("a federated server has no data of its own: workers call federated_update(x, y)")
        if self.mode == "sync":
            while max_versions is None or int(self.engine.version) - start < max_versions:
                if self.step() is None:
                    break
        else:
            def on_step(n: int, loss: torch.Tensor) -> None:
                old = int(self.engine.version) - 1
                self._uploaded(self.queue.ticket, loss)
                self._new_version(old)

            self.engine.run_dataset(self.queue, self.x_all, self.y_all, self.batchSize,
                                    max_steps=max_versions if max_versions is not None else 1 << 30, on_step=on_step)
            self.engine.drain()
            c = self.engine.counters()
            self.numDropped = int(c["dropped"])
            self.done = True
        if self.device.type == "cuda":
            torch.cuda.synchronize()
            self._poll(block=True)
        return {"versions": int(self.engine.version) - start, "uploads": self.numUpdates, "dropped": self.numDropped,
                "loss": None if self.lastLoss is None else float(self.lastLoss), "version": self.model.version}

    def close(self) -> None:
        if self.engine is not None and hasattr(self.engine, "close"):
            self.engine.close()


class B200Worker:
    """This rank's worker handle (`/root/reference/README.md:16-21`, counters of `abstract_client.ts:80-122`)."""

    def __init__(self, server: B200Server, config: Optional[Dict[str, Any]] = None):
        self.server = server
        cfg = dict(config or {})
        if "sendMetrics" in cfg:
            server.sendMetrics = bool(cfg["sendMetrics"])
        if cfg.get("clientId"):
            server.clientId = str(cfg["clientId"])
        self._versions = 0
        self._fx: Optional[torch.Tensor] = None           # federated mode: examples waiting for `examplesPerUpdate`
        self._fy: Optional[torch.Tensor] = None
        server.on_new_version(lambda old, new: setattr(self, "_versions", self._versions + 1))

    @property
    def clientId(self) -> str:
        return self.server.clientId

    @property
    def model(self) -> B200Model:
        return self.server.model

    def setup(self) -> None:
        if self.server.engine is None:
            self.server.setup()

    def model_version(self) -> str:
        return self.server.model.version

    modelVersion = model_version

    # federated mode: the worker's own examples (federated_client.ts:85-128) ----------------------------------------
    def federated_update(self, x: torch.Tensor, y: torch.Tensor) -> int:
        """Add examples; whenever `examplesPerUpdate` have gathered, fit on them and upload (one `federated_upload`).
        Returns the number of uploads made by this call.  Leftover examples wait for the next call."""
        self.setup()
        if x.shape[0] != y.shape[0]:
            raise ValueError("x and y must have the same number of examples")
        self._fx = x if self._fx is None else torch.cat([self._fx, x])
        self._fy = y if self._fy is None else torch.cat([self._fy, y])
        n = int(self.server.clientHyperparams["examplesPerUpdate"])
        made = 0
        while self._fx.shape[0] >= n:
            self.server.federated_upload(self._fx[:n], self._fy[:n])
            self._fx, self._fy = self._fx[n:], self._fy[n:]
            made += 1
        return made

    federatedUpdate = federated_update

    def num_examples(self) -> int:
        return 0 if self._fx is None else int(self._fx.shape[0])

    numExamples = num_examples

    def num_updates(self) -> int:
        return self.server.numUpdates

    def num_versions(self) -> int:
        return self._versions

    numUpdates, numVersions = num_updates, num_versions

    def on_new_version(self, cb) -> None:
        self.server.on_new_version(cb)

    def on_upload(self, cb) -> None:
        self.server.on_upload(cb)

    onNewVersion, onUpload = on_new_version, on_upload

    def run(self, max_versions: Optional[int] = None) -> Dict[str, Any]:
        self.setup()
        return self.server.run(max_versions)

    def evaluate(self, x: torch.Tensor, y: torch.Tensor) -> List[float]:
        """[loss, accuracy] of a batch on this rank's current working weights."""
        s = self.server
        with torch.no_grad():
            out = s.module(x.to(s.device))
            yd = y.to(s.device)
            loss = get_loss(s.compile["loss"])(yd, out)
            lab = yd if yd.dim() == out.dim() - 1 else yd.argmax(-1)
            return [float(loss), float((out.argmax(-1) == lab).float().mean())]
