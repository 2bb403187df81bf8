"""Async / sync servers with real models on CPU: BASELINE config 1 (2 workers on CPU, plumbing), staleness,
re-issue after a lost upload, checkpoint layout + resume, parity with the CPU oracle."""
import json
import os

import numpy as np
import pytest
import torch

from distriflow_b200 import (AsynchronousSGDClient, AsynchronousSGDServer, DistriDataset, DistriServer,
                             DistriServerInMemoryModel, DistriServerTorchModel, DistriWorker, LoopbackHub,
                             SynchronousSGDServer, TcpServerEndpoint, load_layers_model, save_layers_model)
from distriflow_b200.models import mnist_cnn, mnist_mlp, model_from_topology, topology_from_model
from distriflow_b200.oracle import MlpOracle, simulate_async, simulate_sync, synthetic_mnist
from distriflow_b200.transport import AckTimeout


def _data(n=128, seed=0):
    x, y = synthetic_mnist(n, seed=seed)
    return torch.from_numpy(x), torch.from_numpy(y)


def _set_from_oracle(model, oracle):
    with torch.no_grad():
        for p, v in zip(model.parameters(), oracle.vars):
            p.copy_(torch.from_numpy(v))


def test_async_single_worker_matches_oracle():
    torch.manual_seed(0)
    x, y = _data(128)
    oracle = MlpOracle(seed=1, loss="referenceCompat", lr=0.001)
    net = mnist_mlp()
    _set_from_oracle(net, oracle)
    hub = LoopbackHub()
    smodel = DistriServerInMemoryModel(net, {"loss": "referenceCompat", "learningRate": 0.001})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 2})
    server = AsynchronousSGDServer(hub, smodel, ds)
    losses = []
    server.setup()
    wnet = mnist_mlp()
    worker = AsynchronousSGDClient(hub.client, wnet, {"modelCompileConfig": {"loss": "referenceCompat"},
                                                      "sendMetrics": True})
    worker.on_upload(lambda m: losses.append(float(worker.model.lastLoss)))
    worker.setup()
    n = 1 + worker.run()
    assert n == 8 and server.numUpdates == 8 and ds.next().done
    ref = simulate_async(oracle, x.numpy(), y.numpy(), batch_size=32, epochs=2, workers=1)
    assert np.allclose(losses, ref["loss"], rtol=1e-4, atol=1e-5)
    for p, v in zip(net.parameters(), oracle.vars):
        assert np.allclose(p.detach().numpy(), v, rtol=1e-3, atol=1e-5)
    assert worker.lastMetrics is not None and len(worker.lastMetrics) == 2      # [mse, accuracy] pre-step (Q16)


def test_sync_two_workers_cpu_matches_oracle():
    """BASELINE.json config 1: sync batch GD, 2 workers, CPU."""
    x, y = _data(128, seed=3)
    oracle = MlpOracle(seed=2, loss="referenceCompat", lr=0.01)
    net = mnist_mlp()
    _set_from_oracle(net, oracle)
    hub = LoopbackHub()
    smodel = DistriServerInMemoryModel(net, {"loss": "referenceCompat", "learningRate": 0.01})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 1})
    server = DistriServer(hub, smodel, ds, {"serverHyperparams": {"minUpdatesPerVersion": 2}}, mode="sync")
    assert isinstance(server, SynchronousSGDServer)
    versions = []
    server.on_new_version(lambda old, new: versions.append((old, new)))
    server.setup()
    workers = [DistriWorker(hub.client, mnist_mlp(), {"modelCompileConfig": {"loss": "referenceCompat"},
                                                       "clientId": f"w{i}"}, mode="sync") for i in range(2)]
    for w in workers:
        w.setup(process_first=False)
    while any([w.step() for w in workers]):
        pass
    assert len(versions) == 2 and server.workerIDs == {"w0", "w1"}
    ref = simulate_sync(oracle, x.numpy(), y.numpy(), batch_size=32, epochs=1, workers=2)
    assert ref["versions"] == 2
    for p, v in zip(net.parameters(), oracle.vars):
        assert np.allclose(p.detach().numpy(), v, rtol=1e-3, atol=1e-5)


def test_async_staleness_bound_drops_and_reissues():
    x, y = _data(96, seed=5)
    hub = LoopbackHub()
    smodel = DistriServerInMemoryModel(mnist_mlp(), {"loss": "softmaxCrossEntropy", "learningRate": 0.001})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 1})
    server = AsynchronousSGDServer(hub, smodel, ds, {"serverHyperparams": {"maximumStaleness": 0}})
    server.setup()
    w = [AsynchronousSGDClient(hub.client, mnist_mlp(False), {"modelCompileConfig": {"loss": "softmaxCrossEntropy"},
                                                              "clientId": f"w{i}"}) for i in range(2)]
    for c in w:
        c.setup(process_first=False)
    # both hold version v0; w0 uploads first (applied), then w1's gradient is 1 version stale -> dropped
    assert w[0].step() and w[1].step()
    assert server.numDropped == 1 and server.numUpdates == 2
    while any([c.step() for c in w]):
        pass
    assert ds.next().done and ds.epoch == 1
    # every batch was completed exactly by an accepted gradient; the dropped one was re-issued
    assert sum(ds.issued.values()) >= 3 or server.numDropped >= 1


def test_lost_upload_is_reissued():
    x, y = _data(64, seed=7)
    lost = {"n": 0}

    def drop(event, msg):
        if event == "uploadVars" and lost["n"] == 0:
            lost["n"] += 1
            return True
        return False

    hub = LoopbackHub(drop=drop)
    smodel = DistriServerInMemoryModel(mnist_mlp(), {"loss": "referenceCompat"})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 1})
    server = AsynchronousSGDServer(hub, smodel, ds)
    server.setup()
    a = AsynchronousSGDClient(hub.client, mnist_mlp(), {"modelCompileConfig": {"loss": "referenceCompat"}, "clientId": "a"})
    a.setup(process_first=False)
    with pytest.raises(AckTimeout):
        a.step()                                  # worker "dies" after taking batch 0
    b = AsynchronousSGDClient(hub.client, mnist_mlp(), {"modelCompileConfig": {"loss": "referenceCompat"}, "clientId": "b"})
    b.setup(process_first=False)                  # late join gets current weights + next batch
    b.run()
    assert ds.next().done and ds.issued[0] == 2   # batch 0 handed out twice


def test_checkpoint_layout_and_resume(tmp_path):
    net = mnist_mlp()
    m = DistriServerTorchModel(str(tmp_path), net, {"loss": "referenceCompat"})
    m.setup()
    v1 = m.version
    files = sorted(os.listdir(tmp_path / v1))
    assert files == ["model.json", "weights.bin"]
    art = json.load(open(tmp_path / v1 / "model.json"))
    assert art["format"] == "layers-model" and art["weightsManifest"][0]["paths"] == ["weights.bin"]
    names = [w["name"] for w in art["weightsManifest"][0]["weights"]]
    assert names == ["dense_1/kernel", "dense_1/bias", "dense_2/kernel", "dense_2/bias"]
    assert art["weightsManifest"][0]["weights"][0]["shape"] == [784, 10]
    assert os.path.getsize(tmp_path / v1 / "weights.bin") == 7960 * 4
    assert os.path.realpath(tmp_path / "current") == os.path.realpath(tmp_path / v1)
    m.update([torch.ones_like(p) for p in m.get_vars()])
    m.save()
    assert m.list() == [v1, m.version] and os.path.realpath(tmp_path / "current").endswith(m.version)
    # resume picks the newest version and its weights
    m2 = DistriServerTorchModel(str(tmp_path), mnist_mlp, {"loss": "referenceCompat"})
    m2.setup()
    assert m2.version == m.version
    for a, b in zip(m.get_vars(), m2.get_vars()):
        assert torch.equal(a, b)


def test_checkpoint_resume_of_a_non_sequential_model_keeps_batchnorm_statistics(tmp_path):
    """ResNet (not rebuildable from JSON, has BatchNorm buffers): save -> new process state -> `setup()` must come back with the
    trained weights AND the moving statistics, read into the module the caller supplies."""
    from distriflow_b200.models.resnet import resnet20
    torch.manual_seed(0)
    net = resnet20()
    m = DistriServerTorchModel(str(tmp_path), net, {"loss": "softmaxCrossEntropy"})
    m.setup()
    net.train()
    with torch.no_grad():
        net(torch.randn(8, 32, 32, 3))                        # moves every moving_mean / moving_variance
        for p in net.parameters():
            p.add_(0.01)
    m.save(force=True)
    art = json.load(open(tmp_path / m.version / "model.json"))
    names = [w["name"] for w in art["weightsManifest"][0]["weights"]]
    assert any(n.endswith("moving_mean") for n in names) and any(n.endswith("moving_variance") for n in names)
    torch.manual_seed(7)
    m2 = DistriServerTorchModel(str(tmp_path), resnet20(), {"loss": "softmaxCrossEntropy"})
    m2.setup()                                                 # resumes from the newest version directory
    assert m2.version == m.version
    sd1, sd2 = net.state_dict(), m2.model.state_dict()
    assert set(sd1) == set(sd2)
    for k in sd1:
        assert torch.equal(sd1[k], sd2[k]), k
    assert float(sd2["stem.bn.moving_mean"].abs().max()) > 0


def test_reference_cnn_topology_loads_and_round_trips(tmp_path):
    ref = "/root/reference/experiment/mnist/model.json"
    if not os.path.exists(ref):
        pytest.skip("reference checkout not present")
    net = load_layers_model(ref, strict=False)              # shard files are not in the repo
    assert sum(p.numel() for p in net.parameters()) == 600165
    assert net.output_shape == [5] and net.input_shape == [28, 28, 1]
    names = [n for n, _ in net.named_weights()]
    assert names[:2] == ["conv2d_1/kernel", "conv2d_1/bias"] and tuple(net.layers[0].kernel.shape) == (3, 3, 1, 32)
    path = save_layers_model(net, str(tmp_path / "v"))
    back = load_layers_model(path)
    x = torch.rand(2, 28, 28, 1)
    net.eval(), back.eval()
    assert torch.allclose(net(x), back(x), atol=1e-6)
    # multi-group converter-style manifest with separate shard files is readable as well
    art = json.load(open(ref))
    off = 0
    for gi, group in enumerate(art["weightsManifest"]):
        blob = b"".join(dict(net.named_weights())[w["name"]].detach().numpy().tobytes() for w in group["weights"])
        open(tmp_path / group["paths"][0], "wb").write(blob)
    json.dump(art, open(tmp_path / "model.json", "w"))
    again = load_layers_model(str(tmp_path / "model.json"))
    again.eval()
    assert torch.allclose(net(x), again(x), atol=1e-6)


def test_topology_round_trip_matches_builder():
    a = mnist_cnn()
    b = model_from_topology(topology_from_model(a))
    assert [tuple(p.shape) for p in a.parameters()] == [tuple(p.shape) for p in b.parameters()]


def test_async_over_tcp_two_workers():
    import threading
    x, y = _data(128, seed=9)
    ep = TcpServerEndpoint()
    smodel = DistriServerInMemoryModel(mnist_mlp(), {"loss": "referenceCompat"})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 2})
    server = AsynchronousSGDServer(ep, smodel, ds, {"serverHyperparams": {"maximumStaleness": 4}})
    server.setup()
    ws = [AsynchronousSGDClient(ep.url, mnist_mlp(), {"modelCompileConfig": {"loss": "referenceCompat"},
                                                      "clientId": f"w{i}"}) for i in range(2)]
    ths = []
    for w in ws:
        w.setup(process_first=False)
        t = threading.Thread(target=w.run, kwargs={"idle_timeout": 1.0})
        t.start()
        ths.append(t)
    for t in ths:
        t.join(30)
    assert ds.next().done
    assert server.numUpdates - server.numDropped >= 8
    assert server.workerIDs == {"w0", "w1"}
    for w in ws:
        w.dispose()
    ep.close()


def test_worker_crash_mid_batch_is_reissued_to_the_survivor():
    """Fault injection (failure detection): a worker takes a micro-batch and disconnects without uploading.
    The server notices the disconnect (`numClients`), the batch stays in the incomplete set and the dispatcher re-issues it
    to the surviving worker (reference: `dataset.ts:57-60`, `asynchronousSGD_server.ts:54-57`)."""
    x, y = _data(96, seed=11)                       # 3 micro-batches per epoch
    hub = LoopbackHub()
    smodel = DistriServerInMemoryModel(mnist_mlp(), {"loss": "softmaxCrossEntropy", "learningRate": 0.001})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 1})
    server = AsynchronousSGDServer(hub, smodel, ds)
    done_batches = []
    server.on_upload(lambda m: done_batches.append(m.batch))
    server.setup()
    crasher = AsynchronousSGDClient(hub.client, mnist_mlp(), {"clientId": "crasher"})
    crasher.setup(process_first=False)             # connected: holds micro-batch 0, never computes it
    held = crasher.msg.data.batch
    survivor = AsynchronousSGDClient(hub.client, mnist_mlp(), {"clientId": "survivor"})
    survivor.setup(process_first=False)
    assert server.numClients == 2
    crasher.dispose()                              # the crash
    hub.pump()
    assert server.numClients == 1
    n = survivor.run()
    assert ds.next().done, "the epoch must still complete"
    assert sorted(done_batches) == [0, 1, 2] and held in done_batches and n == 3
    assert server.workerIDs == {"survivor"}        # the crasher never delivered anything


@pytest.mark.parametrize("workers,bound,seed", [(2, 0, 1), (3, 1, 2), (4, 2, 3), (3, 1 << 30, 4)])
def test_async_server_invariants_under_random_worker_order(workers, bound, seed):
    """Bounded-staleness async SGD with workers advancing in random order: uploads = applied + dropped, every micro-batch
    of every epoch is completed (re-issued ones possibly twice), no accepted gradient is staler than the bound, versions
    only move forward.  (The reference checks none of this: `asynchronousSGD_server.ts:65-108`)"""
    import random
    rng = random.Random(seed)
    x, y = _data(160, seed=seed)                     # 5 micro-batches per epoch
    hub = LoopbackHub()
    smodel = DistriServerInMemoryModel(mnist_mlp(), {"loss": "softmaxCrossEntropy", "learningRate": 0.001})
    ds = DistriDataset(x, y, {"batchSize": 32, "epochs": 2})
    server = AsynchronousSGDServer(hub, smodel, ds, {"serverHyperparams": {"maximumStaleness": bound}})
    accepted, versions = [], []
    orig_update = server.update_model

    def spy(grads, weight=1.0):
        accepted.append(server.staleness_of(grads.version))
        return orig_update(grads, weight)

    server.update_model = spy
    server.on_new_version(lambda old, new: versions.append((old, new)))
    server.setup()
    ws = [AsynchronousSGDClient(hub.client, mnist_mlp(False), {"modelCompileConfig": {"loss": "softmaxCrossEntropy"},
                                                               "clientId": f"w{i}"}) for i in range(workers)]
    for w in ws:
        w.setup(process_first=False)
    idle = 0
    while idle < 3 * workers:
        w = rng.choice(ws)
        idle = 0 if w.step() else idle + 1
    assert ds.next().done and ds.epoch == 2
    assert server.numUpdates == len(accepted) + server.numDropped
    # every micro-batch of both epochs was completed; a batch re-issued at the end of an epoch while its first holder was
    # still computing yields a second, equally valid gradient, which is applied like the reference does
    # (`asynchronousSGD_server.ts:95-108` applies every upload) — so accepted >= batches, never fewer
    assert 2 * ds.batches <= len(accepted) <= 2 * ds.batches + 2 * (workers - 1)
    assert all(s is not None and 0 <= s <= bound for s in accepted)
    assert len(versions) == len(accepted) and all(o != n for o, n in versions)
    assert [n for _, n in versions[:-1]] == [o for o, _ in versions[1:]]          # a chain: no version skipped or repeated
    if bound >= 1 << 30:
        assert server.numDropped == 0
