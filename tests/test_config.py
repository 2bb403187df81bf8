"""Config override semantics (reference: src/common/utils.ts:157-234)."""
import pytest

from distriflow_b200 import (ConfigError, client_hyperparams, compile_args, dataset_config, engine_config, override,
                             server_hyperparams)


def test_defaults_match_reference():
    assert client_hyperparams() == {"examplesPerUpdate": 5, "learningRate": 0.001, "batchSize": 32, "epochs": 5}
    s = server_hyperparams()
    assert s["aggregation"] == "mean" and s["minUpdatesPerVersion"] == 20
    assert dataset_config() == {"batchSize": 32, "epochs": 5, "smallLastBatch": False}
    assert compile_args() == {"loss": "meanSquaredError", "learningRate": 0.001, "metrics": ["accuracy"]}


def test_override_merges_and_rejects_unknown():
    assert override({"a": 1, "b": 2}, {"b": 3}) == {"a": 1, "b": 3}
    with pytest.raises(ConfigError, match='Unrecognized key "c"'):
        override({"a": 1}, {"c": 2})
    with pytest.raises(ConfigError, match="Error setting clientHyperparams"):
        client_hyperparams({"nope": 1})
    with pytest.raises(ConfigError, match="Error setting serverHyperparams"):
        server_hyperparams({"nope": 1})


def test_zero_is_kept_unlike_reference():
    # the reference's `||` merge turns 0 into the default (utils.ts:210); documented divergence
    assert client_hyperparams({"learningRate": 0})["learningRate"] == 0
    assert client_hyperparams({"learningRate": None})["learningRate"] == 0.001


def test_validation():
    with pytest.raises(ConfigError):
        server_hyperparams({"aggregation": "median"})
    with pytest.raises(ConfigError):
        server_hyperparams({"minUpdatesPerVersion": 0})
    with pytest.raises(ConfigError):
        engine_config({"optimizer": "lamb"})
    assert engine_config({"maximumStaleness": 4})["maximumStaleness"] == 4


def test_packaging_metadata_builds_and_ships_the_native_library():
    """`pip install .` runs the nvcc build (setup.py hook) and the wheel carries the .so (reference: src/*/package.json)."""
    import os
    import tomllib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    meta = tomllib.load(open(os.path.join(root, "pyproject.toml"), "rb"))
    assert meta["project"]["name"] == "distriflow-b200"
    assert "ops/*.so" in meta["tool"]["setuptools"]["package-data"]["distriflow_b200"]
    src = open(os.path.join(root, "setup.py")).read()
    assert "native.build(" in src and "build_py" in src
