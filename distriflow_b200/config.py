"""Hyper-parameter objects and strict override logic.

Capability parity with the reference's config layer
(`/root/reference/src/common/utils.ts:157-234`): typed option bags for the
client, the server, the dataset dispatcher and model compilation, merged with
defaults by a function that rejects unknown keys.  Field names are kept
identical (camelCase) so a DistriFlow user finds every knob where it was.

Deliberate divergences:
  * the reference merges with JavaScript `||`, so an explicit `0`/`false`
    silently falls back to the default (`utils.ts:210`).  Here only `None`
    means "unset".
  * new keys required by the README design / BASELINE north star:
    `maximumStaleness`, `mode`, `optimizer`, `momentum`, `beta1`, `beta2`,
    `eps`, `weightDecay`, `bucketBytes`, `dtype`.
"""
from __future__ import annotations

from typing import Any, Dict, Mapping, Optional

__all__ = [
    "DEFAULT_CLIENT_HYPERPARAMS", "DEFAULT_SERVER_HYPERPARAMS",
    "DEFAULT_DATASET_HYPERPARAMS", "DEFAULT_DISTRIBUTED_COMPILE_ARGS",
    "DEFAULT_ENGINE_CONFIG", "override", "client_hyperparams",
    "server_hyperparams", "dataset_config", "compile_args", "engine_config",
    "ConfigError",
]


class ConfigError(ValueError):
    """Raised for an unknown key or an invalid value in a config bag."""


# reference: utils.ts:181-186
DEFAULT_CLIENT_HYPERPARAMS: Dict[str, Any] = {
    "examplesPerUpdate": 5,
    "learningRate": 0.001,
    "batchSize": 32,
    "epochs": 5,
}

# reference: utils.ts:188-191 (+ maximumStaleness from README.md:27, mode)
DEFAULT_SERVER_HYPERPARAMS: Dict[str, Any] = {
    "aggregation": "mean",
    "minUpdatesPerVersion": 20,
    "maximumStaleness": 0,        # 0 = only gradients of the current version (federated_server.ts:73)
    "stalenessPolicy": "drop",    # "drop" | "scale" (scale = 1/(1+staleness) down-weighting)
}

# reference: utils.ts:193-197
DEFAULT_DATASET_HYPERPARAMS: Dict[str, Any] = {
    "batchSize": 32,
    "epochs": 5,
    "smallLastBatch": False,
}

# reference: utils.ts:199-203
DEFAULT_DISTRIBUTED_COMPILE_ARGS: Dict[str, Any] = {
    "loss": "meanSquaredError",
    "learningRate": 0.001,
    "metrics": ["accuracy"],
}

# no reference counterpart: knobs of the B200 data-parallel engine
DEFAULT_ENGINE_CONFIG: Dict[str, Any] = {
    "mode": "sync",               # "sync" | "async"
    "optimizer": "sgd",           # "sgd" | "momentum" | "adam"
    "learningRate": 0.001,
    "momentum": 0.9,
    "beta1": 0.9,
    "beta2": 0.999,
    "eps": 1e-8,
    "weightDecay": 0.0,
    "maximumStaleness": 4,
    "bucketBytes": 32 << 20,
    "dtype": "bf16",              # working-copy dtype ("bf16" | "fp32")
    "gradDtype": "bf16",          # dtype of the gradient buckets
    "overlap": True,              # launch bucket reduce on a side stream during backward
    "backend": "p2p",             # "p2p" (own kernels) | "nccl" (baseline) | "gloo" (CPU plumbing)
    "fuseAllGather": True,        # all-gather fused into the first consuming GEMM
    "commCtas": 148,              # CTAs of the fused collective kernel (256 threads x <= 128 regs: co-resides with a GEMM CTA)
    "nvls": False,                # gradients reduced and weights broadcast inside the NVSwitch (multicast heap; bf16 only)
    "cudaGraph": False,           # capture the whole step (forward, backward, bucket kernels) into a CUDA graph after warm-up
    "trace": False,               # CUDA-event phase timers + NVTX ranges (forward / backward / collective tail); eager steps only
    "deterministic": False,       # True: reproducible embedding gradient (sorted reduction instead of in-place vector reductions)
    "barrierTimeoutSeconds": 20.0,  # bound of every in-kernel wait for a peer GPU; a waiter that gives up records the silent rank
    "watchdogSeconds": 0.0,       # > 0: host thread reports a rank whose device-written step counter stops advancing
    "reserveSMs": 0,              # SMs the GEMMs leave free during backward (0: rely on co-residency)
    "stalenessPolicy": "drop",    # async: "drop" a gradient older than maximumStaleness, or "scale" it by 1/(1+staleness) up to the bound
    "versionCounts": "updates",   # async: a shard version = one optimizer update (all gradients that met in the accumulation
                                  # buffer are ONE step on their sum) | "gradients": +1 per gradient (reference's literal rule)
    "closeBarrier": "step",       # "step": one closing cross-rank barrier per step; "bucket": one per bucket kernel (round 1)
    "debugPoison": False,         # race detector: NaN a bucket's weights the moment its last gradient is reported (tests)
}

_VALID = {
    "aggregation": {"mean"},                       # federated_server.ts:104 — others throw
    "stalenessPolicy": {"drop", "scale"},
    "mode": {"sync", "async"},
    "optimizer": {"sgd", "momentum", "adam"},
    "dtype": {"bf16", "fp32"},
    "gradDtype": {"bf16", "fp32"},
    "backend": {"p2p", "nccl", "gloo"},
    "closeBarrier": {"step", "bucket"},
    "versionCounts": {"updates", "gradients"},
}


def override(defaults: Mapping[str, Any], choices: Optional[Mapping[str, Any]]) -> Dict[str, Any]:
    """Merge `choices` over `defaults`; unknown keys raise (utils.ts:206-218)."""
    choices = dict(choices or {})
    for key in choices:
        if key not in defaults:
            raise ConfigError(f'Unrecognized key "{key}"')
    out: Dict[str, Any] = {}
    for key, dflt in defaults.items():
        val = choices.get(key)
        out[key] = dflt if val is None else val
        if key in _VALID and out[key] not in _VALID[key]:
            raise ConfigError(f'Invalid value {out[key]!r} for "{key}"; expected one of {sorted(_VALID[key])}')
    return out


def _wrap(name: str, defaults: Mapping[str, Any], hps: Optional[Mapping[str, Any]]) -> Dict[str, Any]:
    try:
        return override(defaults, hps)
    except ConfigError as err:  # same message shape as utils.ts:223,231
        raise ConfigError(f"Error setting {name}: {err}") from None


def client_hyperparams(hps: Optional[Mapping[str, Any]] = None) -> Dict[str, Any]:
    return _wrap("clientHyperparams", DEFAULT_CLIENT_HYPERPARAMS, hps)


def server_hyperparams(hps: Optional[Mapping[str, Any]] = None) -> Dict[str, Any]:
    out = _wrap("serverHyperparams", DEFAULT_SERVER_HYPERPARAMS, hps)
    if int(out["minUpdatesPerVersion"]) < 1:
        raise ConfigError("Error setting serverHyperparams: minUpdatesPerVersion must be >= 1")
    if int(out["maximumStaleness"]) < 0:
        raise ConfigError("Error setting serverHyperparams: maximumStaleness must be >= 0")
    return out


def dataset_config(cfg: Optional[Mapping[str, Any]] = None) -> Dict[str, Any]:
    out = _wrap("datasetConfig", DEFAULT_DATASET_HYPERPARAMS, cfg)
    if int(out["batchSize"]) < 1 or int(out["epochs"]) < 0:
        raise ConfigError("Error setting datasetConfig: batchSize >= 1 and epochs >= 0 required")
    return out


def compile_args(cfg: Optional[Mapping[str, Any]] = None) -> Dict[str, Any]:
    # The reference ctor dereferences an optional config and throws; here None is fine.
    return _wrap("compileArgs", DEFAULT_DISTRIBUTED_COMPILE_ARGS, cfg)


def engine_config(cfg: Optional[Mapping[str, Any]] = None) -> Dict[str, Any]:
    return _wrap("engineConfig", DEFAULT_ENGINE_CONFIG, cfg)
