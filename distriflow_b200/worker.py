"""DistriWorker: gradient computation on dispatched (or own) micro-batches (host protocol layer).

Parity:
  * `AbstractClient` — client id, version tracking, callbacks, counters, upload with
    ack + timeout, connect with timeout (`/root/reference/src/client/abstract_client.ts:12-181`);
  * `AsynchronousSGDClient` — receive (weights, batch) -> optional pre-step metrics ->
    `fit` -> upload `{batch, gradients{version, vars}, clientId, metrics?}`
    (`/root/reference/src/client/asynchronousSGD_client.ts:8-85`);
  * `FederatedClient` — buffers its own examples and uploads one gradient per
    `examplesPerUpdate` (`/root/reference/src/client/federated_client.ts:22-149`);
  * client helpers (`/root/reference/src/client/utils.ts:5-64`): empty-tensor-safe concat/slice,
    `addRows`, cookie helpers (no-ops outside a browser);
  * `DistributedClientModel` flag + `setup()` (`/root/reference/src/client/models.ts:7-34`).

The `server` argument is a URL (`tcp://host:port`) or a callable returning an endpoint
(the reference's `string | () => Socket`, abstract_client.ts:20,166-171).
"""
from __future__ import annotations

import time
import uuid
from typing import Any, Callable, Dict, List, Optional, Sequence, Union

import torch

from .config import DEFAULT_CLIENT_HYPERPARAMS, client_hyperparams
from .messages import DownloadMsg, Events, GradientMsg, UploadMsg
from .model import DistriModel, DistriTorchModel
from .serialization import deserialize_var, serialize_vars
from .transport import AckTimeout, TcpClientEndpoint

__all__ = ["AbstractClient", "AsynchronousSGDClient", "FederatedClient", "DistriWorker",
           "DistriClientTorchModel", "is_distributed_client_model", "add_rows", "concat_with_empty",
           "slice_with_empty", "CONNECTION_TIMEOUT", "UPLOAD_TIMEOUT", "get_cookie", "set_cookie"]

CONNECTION_TIMEOUT = 10.0     # abstract_client.ts:12
UPLOAD_TIMEOUT = 5.0          # abstract_client.ts:13
COOKIE_NAME = "Distributed-learner-uuid"

_CLIENT_CONFIG_KEYS = {"modelCompileConfig", "hyperparams", "verbose", "clientId", "sendMetrics"}
_cookies: Dict[str, str] = {}


def get_cookie(name: str) -> Optional[str]:
    """Outside a browser the reference's cookie helpers do nothing (client/utils.ts:49-64);
    a process-local jar keeps the "same id on reconnect" behaviour."""
    return _cookies.get(name)


def set_cookie(name: str, value: str) -> None:
    _cookies[name] = value


def is_distributed_client_model(model: Any) -> bool:
    return bool(model is not None and (getattr(model, "isDistributedClientModel", False)
                                       or isinstance(model, DistriModel)))


class DistriClientTorchModel(DistriTorchModel):
    isDistributedClientModel = True

    def setup(self) -> None:
        self.fetch_initial()


# ---- empty-tensor-safe helpers (client/utils.ts:22-47) ----------------------------------------

def concat_with_empty(a: Optional[torch.Tensor], b: torch.Tensor) -> torch.Tensor:
    if a is None or a.shape[0] == 0:
        return b.clone()
    return torch.cat([a, b], 0)


def slice_with_empty(t: torch.Tensor, begin: int, size: Optional[int] = None) -> torch.Tensor:
    end = t.shape[0] if size is None else min(begin + size, t.shape[0])
    return t[begin:end]


def add_rows(existing: Optional[torch.Tensor], new: torch.Tensor, input_shape: Sequence[int]) -> torch.Tensor:
    """Append one example (`shape == inputShape`) or a batch (`shape[1:] == inputShape`)."""
    ish = list(input_shape)
    if list(new.shape) == ish:
        return concat_with_empty(existing, new.unsqueeze(0))
    if list(new.shape[1:]) == ish:
        return concat_with_empty(existing, new)
    raise ValueError(f"tensor of shape {list(new.shape)} does not match model shape {ish}")


class AbstractClient:
    def __init__(self, server: Union[str, Callable[[], Any]], model: Any, config: Optional[Dict[str, Any]] = None):
        config = dict(config or {})
        for k in config:
            if k not in _CLIENT_CONFIG_KEYS:
                raise ValueError(f'Unrecognized client config key "{k}"')
        self.server = server
        self.model = model if is_distributed_client_model(model) else \
            DistriClientTorchModel(model, config.get("modelCompileConfig"))
        self.uploadCallbacks: List[Callable[[Optional[UploadMsg]], None]] = []
        self.versionCallbacks: List[Callable[[Optional[str], str], None]] = [
            lambda v1, v2: self.log(f"Updated model: {v1} -> {v2}")]
        self.versionUpdateCounts: Dict[str, int] = {}
        self.verbose = bool(config.get("verbose"))
        self.sendMetrics = bool(config.get("sendMetrics"))
        if config.get("clientId"):
            self.clientId = str(config["clientId"])
        elif get_cookie(COOKIE_NAME):
            self.clientId = get_cookie(COOKIE_NAME)      # type: ignore[assignment]
        else:
            self.clientId = str(uuid.uuid4())
            set_cookie(COOKIE_NAME, self.clientId)
        hp = config.get("hyperparams") or {}
        client_hyperparams(hp)                              # validates keys
        self.hyperparams: Dict[str, Any] = dict(hp)
        self.msg: Optional[DownloadMsg] = None
        self.socket: Any = None
        self.grads: Optional[List[torch.Tensor]] = None
        self.timings: Dict[str, float] = {}

    # -- public surface -----------------------------------------------------------------------
    def model_version(self) -> str:
        return "unsynced" if self.msg is None else self.msg.model.version

    modelVersion = model_version

    def on_new_version(self, cb) -> None:
        self.versionCallbacks.append(cb)

    def on_upload(self, cb) -> None:
        self.uploadCallbacks.append(cb)

    onNewVersion, onUpload = on_new_version, on_upload

    def evaluate(self, x, y) -> List[float]:
        return self.model.evaluate(x, y)

    def predict(self, x):
        return self.model.predict(x)

    def num_updates(self) -> int:
        return sum(self.versionUpdateCounts.values())

    def num_versions(self) -> int:
        return len(self.versionUpdateCounts)

    numUpdates, numVersions = num_updates, num_versions

    def dispose(self) -> None:
        if self.socket is not None:
            self.socket.close()
        self.log("Disconnected")

    @property
    def inputShape(self):
        return self.model.inputShape

    @property
    def outputShape(self):
        return self.model.outputShape

    # -- protected ----------------------------------------------------------------------------
    def log(self, *args: Any) -> None:
        if self.verbose:
            print("Distributed Client:", *args, flush=True)

    def upload_vars(self, msg: UploadMsg) -> Any:
        try:
            return self.socket.emit(Events.Upload, msg, UPLOAD_TIMEOUT)
        except AckTimeout:
            raise AckTimeout("uploadVars timed out") from None

    def set_vars(self, new_vars) -> None:
        self.model.set_vars([deserialize_var(v) for v in new_vars])

    def connect_to(self, server: Union[str, Callable[[], Any]]) -> DownloadMsg:
        self.socket = TcpClientEndpoint(server) if isinstance(server, str) else server()
        self.socket.connect(CONNECTION_TIMEOUT)
        return self.socket.wait_event(Events.Download, CONNECTION_TIMEOUT)

    def time(self, msg: str, action: Callable[[], Any]) -> Any:
        t0 = time.perf_counter()
        out = action()
        dt = (time.perf_counter() - t0) * 1e3
        self.timings[msg] = self.timings.get(msg, 0.0) + dt
        self.log(f"{msg} took {dt:.1f}ms")
        return out

    def _accept_download(self, msg: DownloadMsg, first: bool = False) -> None:
        old = None if first else self.model_version()
        self.msg = msg
        self.set_vars(msg.model.vars)
        new = self.model_version()
        self.versionUpdateCounts.setdefault(new, 0)
        if first or old != new:
            for cb in self.versionCallbacks:
                cb(old, new)


class AsynchronousSGDClient(AbstractClient):
    """Worker for server-owned datasets (async and sync servers speak the same protocol).

    `setup()` connects and performs the first update; later downloads are queued and
    processed by `step()` / `run()` (the reference recurses inside the socket handler,
    asynchronousSGD_client.ts:32-40)."""

    def __init__(self, server, model, config=None):
        super().__init__(server, model, config)
        self._pending: List[DownloadMsg] = []
        self.lastMetrics: Optional[List[float]] = None
        self.batchesDone = 0

    def setup(self, process_first: bool = True) -> None:
        self.time("Initial model setup", self.model.setup)
        first = self.time("Download weights from server", lambda: self.connect_to(self.server))
        self._accept_download(first, first=True)
        self._pending.append(first)
        self.socket.on(Events.Download, self._pending.append)
        if process_first:
            self.step()

    def step(self) -> bool:
        """Process one pending download (newest weights win); False if nothing is pending."""
        if not self._pending:
            if hasattr(self.socket, "hub"):
                self.socket.hub.pump(until=lambda: bool(self._pending))
            if not self._pending:
                return False
        msg = self._pending.pop(0)
        if msg is not self.msg:
            self._accept_download(msg)
        if msg.data is not None:
            self.distributed_update()
        return True

    def run(self, idle_timeout: float = 2.0, max_steps: Optional[int] = None) -> int:
        """Loop until the server stops sending work (dataset exhausted) or `max_steps`."""
        n, idle_since = 0, time.time()
        while max_steps is None or n < max_steps:
            if self.step():
                n += 1
                idle_since = time.time()
            elif hasattr(self.socket, "hub") or time.time() - idle_since > idle_timeout \
                    or not getattr(self.socket, "connected", True):
                break
            else:
                time.sleep(0.001)
        return n

    def distributed_update(self) -> None:
        msg = self.msg
        assert msg is not None and msg.data is not None
        version = self.model_version()
        x, y = deserialize_var(msg.data.x), deserialize_var(msg.data.y)
        metrics = self.model.evaluate(x, y) if self.sendMetrics else None     # pre-step metrics (Q16)
        self.grads = self.time("Fit model", lambda: self.model.fit(x, y))
        grads = serialize_vars(self.grads)
        up = UploadMsg(clientId=self.clientId, batch=msg.data.batch, epoch=msg.data.epoch,
                       gradients=GradientMsg(version=version, vars=grads))
        if metrics is not None:
            up.metrics = metrics
            self.lastMetrics = metrics
        self.time("Upload weights to server", lambda: self.upload_vars(up))
        for cb in self.uploadCallbacks:
            cb(up)
        self.versionUpdateCounts[version] = self.versionUpdateCounts.get(version, 0) + 1
        self.batchesDone += 1

    DistributedUpdate = distributed_update


class FederatedClient(AbstractClient):
    """Worker that trains on its own examples."""

    def __init__(self, server, model, config=None):
        super().__init__(server, model, config)
        self.x: Optional[torch.Tensor] = None
        self.y: Optional[torch.Tensor] = None

    def setup(self) -> None:
        self.time("Initial model setup", self.model.setup)
        first = self.time("Download weights from server", lambda: self.connect_to(self.server))
        self._accept_download(first, first=True)
        self.socket.on(Events.Download, lambda m: self._accept_download(m))

    def distributed_update(self, x: torch.Tensor, y: torch.Tensor) -> None:
        self.x = add_rows(self.x, x, self.model.inputShape)
        self.y = add_rows(self.y, y, self.model.outputShape)
        n = int(self.hyperparam("examplesPerUpdate"))
        while self.x.shape[0] >= n:
            version = self.model_version()
            x_train, y_train = slice_with_empty(self.x, 0, n), slice_with_empty(self.y, 0, n)
            metrics = self.model.evaluate(x_train, y_train) if self.sendMetrics else None
            self.grads = self.time("Fit model", lambda: self.model.fit(x_train, y_train))
            up = UploadMsg(clientId=self.clientId, gradients=GradientMsg(version=version,
                                                                         vars=serialize_vars(self.grads)))
            if metrics is not None:
                up.metrics = metrics
            self.time("Upload weights to server", lambda: self.upload_vars(up))
            for cb in self.uploadCallbacks:
                cb(up)
            self.versionUpdateCounts[version] = self.versionUpdateCounts.get(version, 0) + 1
            self.x, self.y = slice_with_empty(self.x, n), slice_with_empty(self.y, n)

    DistributedUpdate = distributed_update

    def num_examples(self) -> int:
        return 0 if self.x is None else int(self.x.shape[0])

    def hyperparam(self, key: str) -> Any:
        """Precedence local > server-sent > default (federated_client.ts:138-140)."""
        if self.hyperparams.get(key) is not None:
            return self.hyperparams[key]
        if self.msg is not None and self.msg.hyperparams.get(key) is not None:
            return self.msg.hyperparams[key]
        return DEFAULT_CLIENT_HYPERPARAMS[key]

    def num_examples_per_update(self) -> int:
        return int(self.hyperparam("examplesPerUpdate"))

    def num_examples_remaining(self) -> int:
        return self.num_examples_per_update() - self.num_examples()

    numExamples, numExamplesPerUpdate, numExamplesRemaining = num_examples, num_examples_per_update, \
        num_examples_remaining


def DistriWorker(server, model=None, config: Optional[Dict[str, Any]] = None, mode: str = "async", backend: str = "host"):
    """README facade (`/root/reference/README.md:16-21`): 'async'/'sync' -> dispatched batches, 'federated' -> own data.
    `backend="b200"`: `server` is this process's `DistriServer(..., backend="b200")`; the worker is this GPU rank's handle."""
    if backend == "b200":
        from .b200 import B200Worker
        return B200Worker(server, config)
    if mode in ("async", "sync"):
        return AsynchronousSGDClient(server, model, config)
    if mode == "federated":
        return FederatedClient(server, model, config)
    raise ValueError(f"unknown mode {mode}")
