"""DistriServer: parameter server + micro-batch scheduler (host protocol layer).

Parity:
  * `AbstractServer` — state, `onNewVersion` / `onUpload` hooks, `computeDownloadMsg`,
    `log`, `time` (`/root/reference/src/server/abstract_server.ts:24-116`);
  * `AsynchronousSGDServer` — server-owned dataset, every uploaded gradient applied
    at once (`/root/reference/src/server/asynchronousSGD_server.ts:14-110`);
  * `FederatedServer` — buffer K uploads for the current version, mean, one SGD step,
    broadcast (`/root/reference/src/server/federated_server.ts:26-118`);
  * README design: `DistriServer {model, dataset, workerIDs}`
    (`/root/reference/README.md:8-14`).

Changes on purpose: the worker that finished gets the next batch
(Q6) instead of a broadcast of one batch to everybody; gradients carry a version
that is checked against `maximumStaleness` (Q7); a dropped gradient does not
complete its batch, so the dispatcher re-issues it; `SynchronousSGDServer` adds
the README's synchronous batch gradient descent with a server-owned dataset.

This module is the host-side protocol (CPU plumbing config, tests, multi-process
TCP).  On a B200 box the same roles are played by `parallel.engine.DataParallelEngine`,
where the "server" is sharded across the GPU ranks.
"""
from __future__ import annotations

import os
import time
from typing import Any, Callable, Dict, List, Optional, Set


from .checkpoint import DistriServerTorchModel, is_distributed_server_model
from .config import client_hyperparams, server_hyperparams
from .dataset import DistriDataset, batch_to_data_msg
from .messages import DataMsg, DownloadMsg, Events, GradientMsg, ModelMsg, UploadMsg
from .model import StalenessPolicy
from .serialization import deserialize_vars, serialize_vars, stack_serialized

__all__ = ["AbstractServer", "AsynchronousSGDServer", "SynchronousSGDServer", "FederatedServer", "DistriServer"]

_SERVER_CONFIG_KEYS = {"clientHyperparams", "serverHyperparams", "updatesPerVersion", "modelDir",
                       "modelCompileConfig", "verbose", "saveEvery"}


class AbstractServer:
    def __init__(self, endpoint: Any, model: Any, config: Optional[Dict[str, Any]] = None):
        config = dict(config or {})
        for k in config:
            if k not in _SERVER_CONFIG_KEYS:
                raise ValueError(f'Unrecognized server config key "{k}"')
        self.server = endpoint
        self.model = self._wrap_model(model, config)
        self.clientHyperparams = client_hyperparams(config.get("clientHyperparams"))
        self.serverHyperparams = server_hyperparams(config.get("serverHyperparams"))
        self.staleness = StalenessPolicy(self.serverHyperparams["maximumStaleness"],
                                         self.serverHyperparams["stalenessPolicy"])
        self.model.maximumStaleness = self.staleness.maximumStaleness
        self.downloadMsg: Optional[DownloadMsg] = None
        self.numClients = 0
        self.numUpdates = 0
        self.numDropped = 0               # gradients rejected by the staleness rule
        self.updates: List[List[Any]] = []
        self.updating = False
        self.workerIDs: Set[str] = set()  # README.md:12
        self.versionCallbacks: List[Callable[[Optional[str], str], None]] = []
        self.uploadCallbacks: List[Callable[[Optional[UploadMsg]], None]] = []
        verbose = config.get("verbose")
        self.verbose = bool(os.environ.get("VERBOSE")) if verbose is None else bool(verbose)
        self._version_index: Dict[str, int] = {}
        self.timings: Dict[str, float] = {}

    @staticmethod
    def _wrap_model(model: Any, config: Dict[str, Any]) -> Any:
        if is_distributed_server_model(model):
            return model
        model_dir = config.get("modelDir") or os.path.join(os.getcwd(), "saved-models")
        return DistriServerTorchModel(model_dir, model, config.get("modelCompileConfig"),
                                      saveEvery=config.get("saveEvery", 1))

    # hooks (abstract_server.ts:67-79) ---------------------------------------------------------
    def on_new_version(self, cb: Callable[[Optional[str], str], None]) -> None:
        self.versionCallbacks.append(cb)

    def on_upload(self, cb: Callable[[Optional[UploadMsg]], None]) -> None:
        self.uploadCallbacks.append(cb)

    onNewVersion, onUpload = on_new_version, on_upload

    def perform_version_callbacks(self, old_version: Optional[str]) -> None:
        for cb in self.versionCallbacks:
            cb(old_version, self.model.version)

    def perform_upload_callbacks(self, msg: Optional[UploadMsg] = None) -> None:
        for cb in self.uploadCallbacks:
            cb(msg)

    # helpers ---------------------------------------------------------------------------------
    def compute_download_msg(self, data: Optional[DataMsg] = None) -> DownloadMsg:
        """All weights + version + client hyperparams (abstract_server.ts:81-89)."""
        return DownloadMsg(model=ModelMsg(version=self.model.version, vars=serialize_vars(self.model.get_vars())),
                           hyperparams=dict(self.clientHyperparams), data=data)

    computeDownloadMsg = compute_download_msg

    def log(self, *args: Any) -> None:
        if self.verbose:
            print("Distributed Server:", *args, flush=True)

    def time(self, msg: str, action: Callable[[], Any]) -> Any:
        t0 = time.perf_counter()
        out = action()
        dt = (time.perf_counter() - t0) * 1e3
        self.timings[msg] = self.timings.get(msg, 0.0) + dt
        self.log(f"{msg} took {dt:.1f}ms")
        return out

    def _remember_version(self) -> None:
        self._version_index[self.model.version] = self.model.versionIndex
        if len(self._version_index) > 4096:
            for k in list(self._version_index)[:2048]:
                del self._version_index[k]

    def staleness_of(self, grad_version: str) -> Optional[int]:
        idx = self._version_index.get(grad_version)
        return None if idx is None else self.model.versionIndex - idx

    def _track_connection(self, peer: Any) -> None:
        self.numClients += 1
        self.log(f"connection: {self.numClients} clients")

        def on_disconnect(_msg: Any = None, _ack: Any = None) -> None:
            self.numClients -= 1
            self.log(f"disconnection: {self.numClients} clients")

        peer.on(Events.Disconnect, on_disconnect)

    def setup(self) -> None:
        raise NotImplementedError


class AsynchronousSGDServer(AbstractServer):
    """Asynchronous SGD with bounded staleness and a server-owned dataset."""

    def __init__(self, endpoint: Any, model: Any, dataset: DistriDataset, config: Optional[Dict[str, Any]] = None):
        config = dict(config or {})
        # The reference async server never checks the version (asynchronousSGD_server.ts:65-73):
        # keep "apply everything" unless a bound was configured.
        shp = dict(config.get("serverHyperparams") or {})
        shp.setdefault("maximumStaleness", 1 << 30)
        config["serverHyperparams"] = shp
        super().__init__(endpoint, model, config)
        self.dataset = dataset
        self.done = False

    def setup(self) -> None:
        self.time("setting up model", self.model.setup)
        self._remember_version()
        self.server.on_connection(self._on_connection)

    def _send_next(self, peer: Any) -> bool:
        batch = self.dataset.next()
        if batch.done:
            self.done = True
            return False
        self.downloadMsg = self.compute_download_msg(batch_to_data_msg(batch))
        self.log(f"epoch: {batch.epoch} batch: {batch.batch}")
        peer.emit(Events.Download, self.downloadMsg)
        return True

    def _on_connection(self, peer: Any) -> None:
        self._track_connection(peer)
        self._send_next(peer)

        def on_upload(msg: UploadMsg, ack: Callable[[Any], None]) -> None:
            ack(True)
            self.log(f"new update from {msg.clientId}")
            self.workerIDs.add(msg.clientId)
            self.numUpdates += 1
            self.perform_upload_callbacks(msg)
            if msg.gradients is not None:
                s = self.staleness_of(msg.gradients.version)
                w = 0.0 if s is None else self.staleness.weight(s)
                if w > 0.0:
                    self.dataset.complete_batch(msg.batch, msg.epoch)
                    self.update_model(msg.gradients, w)
                else:
                    self.numDropped += 1
                    self.log(f"dropped gradient of version {msg.gradients.version} (staleness {s})")
            self._send_next(peer)

        peer.on(Events.Upload, on_upload)

    def update_model(self, grads: GradientMsg, weight: float = 1.0) -> None:
        self.updating = True
        old = self.model.version
        tensors = deserialize_vars(grads.vars)
        if weight != 1.0:
            tensors = [t * weight for t in tensors]
        self.time("computing new weights", lambda: self.model.update(tensors))
        self.model.save()
        self._remember_version()
        self.updating = False
        self.perform_version_callbacks(old)


class SynchronousSGDServer(AbstractServer):
    """Synchronous batch gradient descent with a server-owned dataset (README.md:1-6).

    Every connected worker gets (weights, micro-batch).  Uploads for the current version
    are summed into `model.savedGradient`; once `minUpdatesPerVersion` arrived the mean is
    applied, the version bumps, and every worker that is waiting receives the new weights
    plus its next micro-batch."""

    def __init__(self, endpoint: Any, model: Any, dataset: DistriDataset, config: Optional[Dict[str, Any]] = None):
        super().__init__(endpoint, model, config)
        self.dataset = dataset
        self.done = False
        self._waiting: List[Any] = []

    def setup(self) -> None:
        self.time("setting up model", self.model.setup)
        self._remember_version()
        self.server.on_connection(self._on_connection)

    def _send_next(self, peer: Any) -> bool:
        batch = self.dataset.next()
        if batch.done:
            self.done = True
            return False
        peer.emit(Events.Download, self.compute_download_msg(batch_to_data_msg(batch)))
        return True

    def _on_connection(self, peer: Any) -> None:
        self._track_connection(peer)
        self._send_next(peer)

        def on_upload(msg: UploadMsg, ack: Callable[[Any], None]) -> None:
            ack(True)
            self.workerIDs.add(msg.clientId)
            self.perform_upload_callbacks(msg)
            s = self.staleness_of(msg.gradients.version) if msg.gradients is not None else None
            w = 0.0 if s is None else self.staleness.weight(s)
            if w <= 0.0:
                self.numDropped += 1
                self._send_next(peer)            # fresh weights, fresh batch; the old batch gets re-issued
                return
            self.numUpdates += 1
            self.dataset.complete_batch(msg.batch, msg.epoch)
            self.model.accumulate(deserialize_vars(msg.gradients.vars), w)
            self._waiting.append(peer)
            if self.model.savedCount >= int(self.serverHyperparams["minUpdatesPerVersion"]):
                self._step()

        peer.on(Events.Upload, on_upload)

    def _step(self) -> None:
        self.updating = True
        old = self.model.version
        self.time("computing new weights", lambda: self.model.apply_saved(self.serverHyperparams["aggregation"]))
        self.model.save()
        self._remember_version()
        self.updating = False
        self.perform_version_callbacks(old)
        waiting, self._waiting = self._waiting, []
        for p in waiting:
            self._send_next(p)

    def flush(self) -> None:
        """Apply a partial buffer (fewer than `minUpdatesPerVersion` contributions), e.g. at epoch end."""
        if self.model.savedCount:
            self._step()


class FederatedServer(AbstractServer):
    """Workers own their data; the server averages K gradient uploads per version."""

    def setup(self) -> None:
        self.time("setting up model", self.model.setup)
        self._remember_version()
        self.downloadMsg = self.compute_download_msg()
        self.perform_version_callbacks(None)
        self.server.on_connection(self._on_connection)

    def _on_connection(self, peer: Any) -> None:
        self._track_connection(peer)
        peer.emit(Events.Download, self.downloadMsg)

        def on_upload(msg: UploadMsg, ack: Callable[[Any], None]) -> None:
            ack(True)
            self.workerIDs.add(msg.clientId)
            s = self.staleness_of(msg.gradients.version) if msg.gradients is not None else None
            ok = s is not None and self.staleness.weight(s) > 0.0 and not self.updating
            if ok:                                  # federated_server.ts:73
                self.log(f"new update from {msg.clientId}")
                self.updates.append(msg.gradients.vars)
                self.numUpdates += 1
                self.perform_upload_callbacks(msg)
                if self.should_update():
                    self.update_model()
                    self.server.broadcast(Events.Download, self.downloadMsg)
            else:
                self.numDropped += 1

        peer.on(Events.Upload, on_upload)

    def should_update(self) -> bool:
        return self.numUpdates >= int(self.serverHyperparams["minUpdatesPerVersion"])

    def update_model(self) -> None:
        self.updating = True
        old = self.model.version
        if self.serverHyperparams["aggregation"] != "mean":
            raise ValueError(f"unsupported aggregation {self.serverHyperparams['aggregation']}")

        def compute() -> None:
            stacked = deserialize_vars(stack_serialized(self.updates))    # [K, ...] per weight
            self.model.update([t.float().mean(0) for t in stacked])

        self.time("computing new weights", compute)
        self.model.save()
        self._remember_version()
        self.downloadMsg = self.compute_download_msg()
        self.updates = []
        self.numUpdates = 0
        self.updating = False
        self.perform_version_callbacks(old)


def DistriServer(endpoint: Any, model: Any, dataset: Optional[DistriDataset] = None,
                 config: Optional[Dict[str, Any]] = None, mode: str = "async", backend: str = "host", **kw: Any):
    """README facade: `mode` in {'async', 'sync', 'federated'}.  `backend="host"`: the message protocol over `endpoint`
    (in-process hub or TCP).  `backend="b200"`: every rank of the job builds the same server, which is then sharded across
    the GPU ranks (`b200.B200Server`; `endpoint` is unused and may be None; 'federated': no dataset, workers bring their examples)."""
    if backend == "b200":
        from .b200 import B200Server
        return B200Server(model, dataset, config, mode=mode, **kw)
    if backend != "host":
        raise ValueError(f"unknown backend {backend}")
    if mode == "async":
        return AsynchronousSGDServer(endpoint, model, dataset, config)
    if mode == "sync":
        return SynchronousSGDServer(endpoint, model, dataset, config)
    if mode == "federated":
        return FederatedServer(endpoint, model, config)
    raise ValueError(f"unknown mode {mode}")
