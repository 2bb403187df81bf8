"""Checkpoint IO in the tf.js LayersModel layout + the server-side model variants.

Parity:
  * `model.save('file://<dir>/<version>')` / `tf.loadLayersModel('<dir>/<version>/model.json')`
    as used by `DistributedServerTfModel` (`/root/reference/src/server/models.ts:98-150`):
    `<saveDir>/<ms-timestamp>/{model.json,weights.bin}`, a `current` symlink,
    auto-resume from the newest version on `setup()`;
  * `DistributedServerInMemoryModel` (`models.ts:63-75`) — versions, no files;
  * `DistributedServerDynamicModel` (`models.ts:153-218`) — `meta.json` + `data.bin`;
  * `fetchModel` (`/root/reference/src/common/utils.ts:236-244`).

The reference cannot actually write `file://` (needs tfjs-node), so the
layout is written here directly.  Reference bugs fixed: `list()` drops only
`current` (Q10); dynamic `load()` assigns what it read and `save()` creates the
directory (Q11); a checkpoint is written on an interval instead of on every
applied gradient (Q8) — the version counter still bumps on every update.

Accepted on load: one manifest group with `weights.bin`, or the converter style
of `/root/reference/experiment/mnist/model.json` (several groups, each with its
own shard paths, topology wrapped with `keras_version`/`model_config`).
"""
from __future__ import annotations

import json
import os
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .model import DistriDynamicModel, DistriTorchModel
from .serialization import DTYPES, flat_deserialize, flat_serialize

__all__ = ["save_layers_model", "load_layers_model", "fetch_model", "force_symlink", "new_version",
           "DistriServerInMemoryModel", "DistriServerTorchModel", "DistriServerDynamicModel",
           "is_distributed_server_model", "DistributedServerInMemoryModel", "DistributedServerTfModel",
           "DistributedServerDynamicModel", "isDistributedServerModel", "list_versions"]

_last_version = 0


def new_version() -> str:
    """Decimal milliseconds since epoch, strictly increasing within the process."""
    global _last_version
    v = int(time.time() * 1000)
    if v <= _last_version:
        v = _last_version + 1
    _last_version = v
    return str(v)


def force_symlink(src: str, dest: str) -> None:
    """`dest -> src`, replacing a stale link (models.ts:17-30)."""
    try:
        os.symlink(src, dest, target_is_directory=True)
    except FileExistsError:
        if os.path.islink(dest) and os.readlink(dest) == src:
            return
        tmp = f"{dest}.tmp{os.getpid()}"
        os.symlink(src, tmp, target_is_directory=True)
        os.replace(tmp, dest)


# ----------------------------------------------------------------------------------------------
# tf.js layers-model artifact

def _named_weights(model: torch.nn.Module) -> List[Tuple[str, torch.Tensor]]:
    """Everything a restart needs: trainable weights AND persistent buffers (BatchNorm moving_mean / moving_variance),
    which Keras / tf.js list in the manifest as non-trainable weights of the layer."""
    if hasattr(model, "named_weights"):
        return model.named_weights()
    out = [(n.replace(".", "/"), p) for n, p in model.named_parameters()]
    nonpersistent = set()
    for mn, m in model.named_modules():
        for bn in getattr(m, "_non_persistent_buffers_set", ()):
            nonpersistent.add(f"{mn}.{bn}" if mn else bn)
    out += [(n.replace(".", "/"), b) for n, b in model.named_buffers() if n not in nonpersistent and b.is_floating_point()]
    return out


def save_layers_model(model: torch.nn.Module, directory: str, topology: Optional[Dict[str, Any]] = None) -> str:
    """Write `<directory>/model.json` + `<directory>/weights.bin`; returns the model.json path."""
    os.makedirs(directory, exist_ok=True)
    if topology is None:
        from .models.topology import SequentialModel, topology_from_model
        if isinstance(model, SequentialModel):
            topology = topology_from_model(model)
        else:                                       # not rebuildable from JSON: class name + its config for the record
            import dataclasses
            cfg = getattr(model, "config", {})
            if dataclasses.is_dataclass(cfg) and not isinstance(cfg, type):
                cfg = dataclasses.asdict(cfg)
            topology = {"class_name": model.__class__.__name__, "config": cfg if isinstance(cfg, dict) else {}}
    specs, blobs = [], []
    for name, t in _named_weights(model):
        t32 = t.detach().to("cpu", torch.float32).contiguous()     # weights.bin is float32
        specs.append({"name": name, "shape": list(t32.shape), "dtype": "float32"})
        blobs.append(t32.numpy().tobytes())
    artifact = {"modelTopology": topology, "format": "layers-model",
                "generatedBy": "distriflow_b200", "convertedBy": None,
                "weightsManifest": [{"paths": ["weights.bin"], "weights": specs}]}
    tmp = os.path.join(directory, "weights.bin.tmp")
    with open(tmp, "wb") as f:
        for b in blobs:
            f.write(b)
    os.replace(tmp, os.path.join(directory, "weights.bin"))
    path = os.path.join(directory, "model.json")
    with open(path + ".tmp", "w") as f:
        json.dump(artifact, f)
    os.replace(path + ".tmp", path)
    return path


def read_manifest_weights(model_json: str, artifact: Dict[str, Any]) -> Dict[str, torch.Tensor]:
    base = os.path.dirname(os.path.abspath(model_json))
    out: Dict[str, torch.Tensor] = {}
    for group in artifact.get("weightsManifest", []):
        blob = b"".join(open(os.path.join(base, p), "rb").read() for p in group["paths"])
        off = 0
        for w in group["weights"]:
            np_dtype = np.dtype(DTYPES[w["dtype"]][0])
            n = int(np.prod(w["shape"])) if w["shape"] else 1
            nbytes = n * np_dtype.itemsize
            if off + nbytes > len(blob):
                raise ValueError(f"weight {w['name']} runs past the end of {group['paths']}")
            arr = np.frombuffer(blob, dtype=np_dtype, count=n, offset=off).copy()
            out[w["name"]] = torch.from_numpy(arr).reshape(w["shape"])
            off += nbytes
    return out


def load_layers_model(path: str, builder: Optional[Callable[[Dict[str, Any]], torch.nn.Module]] = None,
                      load_weights: bool = True, strict: bool = True, into: Optional[torch.nn.Module] = None) -> torch.nn.Module:
    """`path` is a model.json or a directory holding one.  `into`: read the manifest into this existing module instead of
    rebuilding one from the topology (needed for models that are not Sequential: ResNet, GPT-2, user modules)."""
    if os.path.isdir(path):
        path = os.path.join(path, "model.json")
    with open(path) as f:
        artifact = json.load(f)
    topo = artifact["modelTopology"]
    if into is not None:
        model = into
    else:
        if builder is None:
            from .models.topology import model_from_topology
            builder = model_from_topology
        model = builder(topo)
    if load_weights:
        try:
            weights = read_manifest_weights(path, artifact)
        except FileNotFoundError:
            if strict:
                raise
            weights = {}        # e.g. the reference repo ships model.json without its shards
        named = dict(_named_weights(model))
        missing = [n for n in named if n not in weights]
        if weights and missing and strict:
            raise KeyError(f"weights missing from manifest: {missing}")
        with torch.no_grad():
            for n, t in named.items():
                if n in weights:
                    if tuple(weights[n].shape) != tuple(t.shape):
                        raise ValueError(f"{n}: manifest shape {tuple(weights[n].shape)} != model {tuple(t.shape)}")
                    t.copy_(weights[n].to(t.dtype))
    return model


def fetch_model(async_model: Any) -> torch.nn.Module:
    """str -> load from path; callable -> call it; module -> pass through (utils.ts:236-244)."""
    if isinstance(async_model, str):
        return load_layers_model(async_model)
    if isinstance(async_model, torch.nn.Module):
        return async_model
    if callable(async_model):
        return async_model()
    raise TypeError(f"cannot resolve a model from {type(async_model)}")


def list_versions(save_dir: str) -> List[str]:
    """Version directories, oldest first; only the `current` link is filtered out (Q10 fixed)."""
    if not os.path.isdir(save_dir):
        return []
    names = [n for n in os.listdir(save_dir) if n != "current" and os.path.isdir(os.path.join(save_dir, n))
             and not n.endswith(".tmp")]
    return sorted(names, key=lambda s: (len(s), s))


# ----------------------------------------------------------------------------------------------
# server model variants

def is_distributed_server_model(model: Any) -> bool:
    return bool(model is not None and getattr(model, "isDistributedServerModel", False))


isDistributedServerModel = is_distributed_server_model


class DistriServerInMemoryModel(DistriTorchModel):
    """Versions without files (models.ts:63-75)."""
    isDistributedServerModel = True

    def setup(self) -> None:
        self.fetch_initial()
        self.save()

    def save(self) -> None:
        self.version = new_version()
        self.versionIndex += 1


class DistriServerTorchModel(DistriTorchModel):
    """File-backed server model: version directories under `saveDir` (models.ts:85-151)."""
    isDistributedServerModel = True

    def __init__(self, saveDir: str, initialModel: Any = None, config: Optional[Dict[str, Any]] = None,
                 saveEvery: int = 1, device: Optional[torch.device] = None, keepLast: Optional[int] = None):
        super().__init__(initialModel, config, device)
        self.saveDir = saveDir
        self.saveEvery = max(1, int(saveEvery))     # checkpoint interval in versions (Q8)
        self.keepLast = keepLast
        self._since_save = 0

    def setup(self) -> None:
        os.makedirs(self.saveDir, exist_ok=True)
        last = self.last()
        if last:
            self.load(last)
        else:
            self.fetch_initial()
            self.save(force=True)

    def list(self) -> List[str]:
        return list_versions(self.saveDir)

    def last(self) -> Optional[str]:
        models = self.list()
        return models[-1] if models else None

    def save(self, force: bool = False) -> None:
        self.version = new_version()
        self.versionIndex += 1
        self._since_save += 1
        if not force and self._since_save < self.saveEvery:
            return
        self._since_save = 0
        final = os.path.join(self.saveDir, self.version)
        save_layers_model(self.model, final + ".tmp")
        os.replace(final + ".tmp", final)
        force_symlink(final, os.path.join(self.saveDir, "current"))
        if self.keepLast:
            import shutil
            for old in self.list()[:-self.keepLast]:
                shutil.rmtree(os.path.join(self.saveDir, old), ignore_errors=True)

    def load(self, version: str) -> None:
        path = os.path.join(self.saveDir, version, "model.json")
        self.version = version
        with open(path) as f:
            topo = json.load(f)["modelTopology"]
        from .models.topology import is_sequential_topology
        if is_sequential_topology(topo):
            self.model = load_layers_model(path)
        else:
            # ResNet / GPT-2 / user modules are not rebuilt from JSON: the weights are read into the module the caller
            # supplied (the reference can always rebuild because tf.js LayersModels are self-describing, models.ts:140-150)
            self._resolve()
            if self.model is None:
                raise ValueError(f"{path}: topology {topo.get('class_name')!r} cannot be rebuilt; pass the module as initialModel")
            load_layers_model(path, into=self.model)
        if self.device is not None:
            self.model.to(self.device)
        self._params = [p for p in self.model.parameters() if p.requires_grad]
        force_symlink(os.path.join(self.saveDir, version), os.path.join(self.saveDir, "current"))


class DistriServerDynamicModel(DistriDynamicModel):
    """Functional server model persisted as `<saveDir>/<version>/{meta.json,data.bin}` (models.ts:153-218)."""
    isDistributedServerModel = True

    def __init__(self, *, saveDir: str, vars: Sequence[torch.Tensor], predict: Callable, loss: Callable,
                 inputShape: Sequence[int], outputShape: Sequence[int], learningRate: float = 0.001):
        super().__init__(vars=vars, predict=predict, loss=loss, inputShape=inputShape, outputShape=outputShape,
                         learningRate=learningRate)
        self.saveDir = saveDir
        self.version = ""

    def setup(self) -> None:
        os.makedirs(self.saveDir, exist_ok=True)
        last = self.last()
        if last:
            self.load(last)
        else:
            self.save()

    def list(self) -> List[str]:
        return list_versions(self.saveDir)

    def last(self) -> Optional[str]:
        models = self.list()
        return models[-1] if models else None

    def save(self) -> None:
        self.version = new_version()
        self.versionIndex += 1
        d = os.path.join(self.saveDir, self.version)
        os.makedirs(d + ".tmp", exist_ok=True)
        data, meta = flat_serialize(self.get_vars())
        with open(os.path.join(d + ".tmp", "data.bin"), "wb") as f:
            f.write(data)
        with open(os.path.join(d + ".tmp", "meta.json"), "w") as f:
            json.dump(meta, f)
        os.replace(d + ".tmp", d)
        force_symlink(d, os.path.join(self.saveDir, "current"))

    def load(self, version: str) -> None:
        d = os.path.join(self.saveDir, version)
        with open(os.path.join(d, "meta.json")) as f:
            meta = json.load(f)
        with open(os.path.join(d, "data.bin"), "rb") as f:
            data = f.read()
        self.set_vars(flat_deserialize(data, meta))      # Q11: assign what was read
        self.version = version
        force_symlink(d, os.path.join(self.saveDir, "current"))


DistributedServerInMemoryModel = DistriServerInMemoryModel
DistributedServerTfModel = DistriServerTorchModel
DistributedServerDynamicModel = DistriServerDynamicModel
