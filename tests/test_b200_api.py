"""`DistriServer(..., backend="b200")`: the reference's API scenarios on the engine path.

The three scenarios of `/root/reference/src/test/federated_api_test.ts:56-90` (version visible on start-up, an update is
transmitted, a new version after enough uploads with `onNewVersion(old, new)`), plus what the README design adds
(`workerIDs`, `savedGradient`, `sendMetrics`, `saveEvery`).  Here on the CPU back end of the engine (one rank); the same
file runs them on a B200 with the device work queue and the metric ring (`-m gpu`)."""
import json
import os

import pytest
import torch

from distriflow_b200 import DistriDataset, DistriServer, DistriWorker
from distriflow_b200.models import mnist_mlp
from distriflow_b200.oracle import synthetic_mnist


def _make(device, n=64, cfg=None, ds_cfg=None, seed=0):
    x, y = synthetic_mnist(n, seed=1)
    ds = DistriDataset(torch.from_numpy(x) / 255.0, torch.from_numpy(y), {"batchSize": 8, "epochs": 1, **(ds_cfg or {})})
    torch.manual_seed(seed)
    conf = {"modelCompileConfig": {"loss": "softmaxCrossEntropy", "learningRate": 0.05}, **(cfg or {})}
    server = DistriServer(None, mnist_mlp(softmax_output=False), ds, conf, mode="sync", backend="b200", device=device)
    return server, DistriWorker(server, backend="b200")


DEVICES = [pytest.param(torch.device("cpu"), id="cpu"),
           pytest.param(torch.device("cuda", 0), id="b200", marks=pytest.mark.gpu)]


@pytest.mark.parametrize("device", DEVICES)
def test_transmits_model_version_on_startup(device):
    server, worker = _make(device)
    worker.setup()
    assert worker.model_version() == server.model.version == "0"
    assert server.workerIDs == {worker.clientId} and server.numClients == 1


@pytest.mark.parametrize("device", DEVICES)
def test_transmits_updates_and_new_version_after_enough_uploads(device):
    server, worker = _make(device, cfg={"serverHyperparams": {"minUpdatesPerVersion": 2}})
    seen, uploads = [], []
    worker.on_new_version(lambda old, new: seen.append((old, new)))
    server.on_upload(uploads.append)
    worker.setup()
    assert server.upload() is not None
    assert server.numUpdates == 1 and server.model.savedCount == 1 and server.model.version == "0" and seen == []
    assert server.model.savedGradient is not None                 # the aggregated gradient lives in the engine's buckets
    assert server.upload() is not None
    assert server.model.version == "1" and seen == [("0", "1")] and server.model.savedCount == 0
    assert worker.num_updates() == 2 and worker.num_versions() == 1
    out = server.run()
    assert out["uploads"] == 8 and server.model.version == "4" and server.done
    assert [u.batch for u in uploads] == list(range(8)) and all(u.clientId == worker.clientId for u in uploads)


@pytest.mark.parametrize("device", DEVICES)
def test_accumulated_micro_batches_equal_one_large_batch(device):
    """K micro-batches per version = the mean gradient of the K batches = one SGD step on their union."""
    a, _ = _make(device, cfg={"serverHyperparams": {"minUpdatesPerVersion": 4}})
    b, _ = _make(device, cfg={"serverHyperparams": {"minUpdatesPerVersion": 1}}, ds_cfg={"batchSize": 32})
    a.run()
    b.run()
    assert a.model.version == b.model.version == "2"
    for wa, wb in zip(a.model.get_vars(), b.model.get_vars()):
        assert torch.allclose(wa.cpu(), wb.cpu(), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("device", DEVICES)
def test_send_metrics_and_checkpoint_cadence(device, tmp_path):
    server, worker = _make(device, n=96, cfg={"sendMetrics": True, "modelDir": str(tmp_path), "saveEvery": 5,
                                               "serverHyperparams": {"minUpdatesPerVersion": 1}})
    got = []
    server.on_upload(got.append)
    out = worker.run()
    assert out["versions"] == 12 and len(got) == 12
    assert all(m.metrics is not None and len(m.metrics) == 2 and 0.0 <= m.metrics[1] <= 1.0 for m in got)
    assert got[-1].metrics[0] < got[0].metrics[0]                # the batch loss falls
    versions = sorted(d for d in os.listdir(tmp_path) if d != "current")
    assert len(versions) == 2                                    # versions 5 and 10
    art = json.load(open(tmp_path / "current" / "model.json"))
    assert [w["name"] for w in art["weightsManifest"][0]["weights"]][:2] == ["dense_1/kernel", "dense_1/bias"]
    loss, acc = worker.evaluate(server.dataset.x[:32], server.dataset.y[:32])
    assert loss == loss and 0.0 <= acc <= 1.0


def test_unknown_backend_and_keys_are_rejected():
    x, y = synthetic_mnist(16, seed=1)
    ds = DistriDataset(torch.from_numpy(x), torch.from_numpy(y), {"batchSize": 8, "epochs": 1})
    with pytest.raises(ValueError):
        DistriServer(None, mnist_mlp(), ds, {}, mode="sync", backend="tpu")
    with pytest.raises(ValueError):
        DistriServer(None, mnist_mlp(), ds, {"bogus": 1}, mode="sync", backend="b200")
    with pytest.raises(ValueError):
        DistriServer(None, mnist_mlp(), None, {}, mode="sync", backend="b200")


def test_federated_mode_workers_bring_their_own_examples():
    """`mode="federated"` on the engine path (CPU back end here): no server-owned dataset; the worker buffers its own examples
    until `examplesPerUpdate`, every full buffer is one upload, `minUpdatesPerVersion` uploads make a version
    (/root/reference/src/client/federated_client.ts:85-128, /root/reference/src/server/federated_server.ts:71-116)."""
    x, y = synthetic_mnist(40, seed=2)
    x, y = torch.from_numpy(x) / 255.0, torch.from_numpy(y)
    torch.manual_seed(0)
    server = DistriServer(None, mnist_mlp(softmax_output=False), None,
                          {"modelCompileConfig": {"loss": "softmaxCrossEntropy", "learningRate": 0.05},
                           "clientHyperparams": {"examplesPerUpdate": 8},
                           "serverHyperparams": {"minUpdatesPerVersion": 2}, "sendMetrics": True},
                          mode="federated", backend="b200", device=torch.device("cpu"))
    worker = DistriWorker(server, backend="b200")
    versions, uploads = [], []
    server.on_new_version(lambda old, new: versions.append((old, new)))
    server.on_upload(lambda m: uploads.append(m))
    assert worker.federated_update(x[:5], y[:5]) == 0 and worker.num_examples() == 5          # not enough yet
    assert worker.federated_update(x[5:20], y[5:20]) == 2 and worker.num_examples() == 4     # 20 examples: two uploads, 4 left
    assert versions == [("0", "1")] and len(uploads) == 2 and uploads[0].metrics is not None
    assert worker.federated_update(x[20:40], y[20:40]) == 3 and worker.num_examples() == 0
    assert [v[1] for v in versions] == ["1", "2"] and server.model.savedCount == 1            # the fifth upload waits for a sixth
    with pytest.raises(RuntimeError):
        server.run()
    # a version is the MEAN of its uploads: same result as one step on the 16 examples of the two buffers
    torch.manual_seed(0)
    from distriflow_b200.losses import get_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    ref = mnist_mlp(softmax_output=False)
    ce = get_loss("softmaxCrossEntropy")
    eng = DataParallelEngine(ref, lambda m, a, b: ce(b, m(a)), {"optimizer": "sgd", "learningRate": 0.05, "dtype": "fp32",
                                                               "gradDtype": "fp32"}, device=torch.device("cpu"))
    eng.step(x[:16], y[:16])
    eng.step(x[16:32], y[16:32])
    for a, b in zip(server.model.get_vars(), eng.full_precision_weights()):
        assert torch.allclose(a, b, atol=1e-6)
