"""Model wrappers, server model variants, client helpers, transport timeouts (reference: src/common/models.ts,
src/server/models.ts, src/client/utils.ts, src/client/abstract_client.ts)."""
import json
import os

import pytest
import torch

from distriflow_b200 import (DistriDynamicModel, DistriServerDynamicModel, DistriServerInMemoryModel, DistriTorchModel,
                             FederatedClient, FederatedServer, LoopbackHub, StalenessPolicy, is_distributed_server_model)
from distriflow_b200.checkpoint import force_symlink, list_versions, new_version
from distriflow_b200.losses import LOSSES, get_loss
from distriflow_b200.models import mnist_mlp
from distriflow_b200.transport import AckTimeout, ConnectTimeout, TcpClientEndpoint
from distriflow_b200.worker import add_rows, concat_with_empty, get_cookie, set_cookie, slice_with_empty


def test_fit_returns_gradients_and_leaves_weights_untouched():
    torch.manual_seed(0)
    m = DistriTorchModel(mnist_mlp(), {"loss": "categoricalCrossentropy", "learningRate": 0.1})
    before = [v.clone() for v in m.get_vars()]
    x, y = torch.rand(8, 28, 28, 1), torch.eye(10)[torch.randint(0, 10, (8,))]
    grads = m.fit(x, y)
    assert [tuple(g.shape) for g in grads] == [(784, 10), (10,), (10, 10), (10,)]      # parameter-list order (Q14)
    assert all(torch.equal(a, b) for a, b in zip(before, m.get_vars()))
    m.update(grads)
    assert all(torch.allclose(b - 0.1 * g, a) for a, b, g in zip(m.get_vars(), before, grads))
    ev = m.evaluate(x, y)
    assert len(ev) == 2 and 0.0 <= ev[1] <= 1.0                                          # [loss, accuracy]
    assert m.inputShape == [28, 28, 1] and m.outputShape == [10]
    with pytest.raises(ValueError):
        m.set_vars(before[:2])


def test_config_is_optional_and_loss_names_of_the_reference_exist():
    DistriTorchModel(mnist_mlp())                                                       # reference ctor throws here (Q5)
    for name in ("absoluteDifference", "computeWeightedLoss", "hingeLoss", "huberLoss", "logLoss", "meanSquaredError",
                 "sigmoidCrossEntropy", "softmaxCrossEntropy"):
        assert name in LOSSES
        v = get_loss(name)(torch.rand(4, 3), torch.rand(4, 3))
        assert v.dim() == 0 and torch.isfinite(v)
    with pytest.raises(KeyError):
        get_loss("cosineDistance")                                                      # commented out in the reference too


def test_dynamic_model_update_is_sgd_not_nan():
    w = torch.ones(3, 2)
    m = DistriDynamicModel(vars=[w], predict=lambda x: x @ w, loss=lambda y, p: ((y - p) ** 2).mean(-1),
                           inputShape=[3], outputShape=[2], learningRate=0.5)
    x, y = torch.ones(4, 3), torch.zeros(4, 2)
    g = m.fit(x, y)
    m.update(g)
    assert torch.isfinite(m.get_vars()[0]).all() and torch.allclose(m.get_vars()[0], torch.ones(3, 2) - 0.5 * g[0])   # Q12
    assert len(m.evaluate(x, y)) == 4


def test_server_dynamic_model_saves_and_resumes(tmp_path):
    def make(v):
        return DistriServerDynamicModel(saveDir=str(tmp_path / "dyn"), vars=[v], predict=lambda x: x, loss=lambda y, p: (y - p) ** 2,
                                        inputShape=[2], outputShape=[2])
    a = make(torch.tensor([1.0, 2.0]))
    a.setup()                                                                           # reference ctor saves before mkdir (Q11)
    v1 = a.version
    assert sorted(os.listdir(tmp_path / "dyn" / v1)) == ["data.bin", "meta.json"]
    meta = json.load(open(tmp_path / "dyn" / v1 / "meta.json"))
    assert meta["byteOffsets"] == [0] and meta["meta"][0]["shape"] == [2]
    a.update([torch.tensor([1.0, 1.0])])
    a.save()
    b = make(torch.zeros(2))
    b.setup()
    assert b.version == a.version and torch.allclose(b.get_vars()[0], a.get_vars()[0])   # load really assigns (Q11)
    assert is_distributed_server_model(b) and not is_distributed_server_model(object())


def test_version_listing_and_symlink(tmp_path):
    for name in ("100", "99", "1000"):
        os.makedirs(tmp_path / name)
    force_symlink(str(tmp_path / "99"), str(tmp_path / "current"))
    assert list_versions(str(tmp_path)) == ["99", "100", "1000"]                         # `current` filtered, nothing else (Q10)
    force_symlink(str(tmp_path / "1000"), str(tmp_path / "current"))
    assert os.path.realpath(tmp_path / "current").endswith("1000")
    a, b = new_version(), new_version()
    assert int(b) > int(a)


def test_in_memory_model_versions_and_saved_gradient():
    m = DistriServerInMemoryModel(mnist_mlp(), {"loss": "meanSquaredError", "learningRate": 1.0})
    m.setup()
    v0, before = m.version, [v.clone() for v in m.get_vars()]
    g = [torch.ones_like(v) for v in before]
    assert m.accumulate(g) == 1 and m.accumulate([2 * t for t in g]) == 2
    assert m.apply_saved("mean") == 2 and m.savedGradient is None                        # README savedGradient
    assert all(torch.allclose(b - 1.5, a) for a, b in zip(m.get_vars(), before))
    m.save()
    assert m.version != v0 and m.modelID == m.version
    with pytest.raises(ValueError):
        m.accumulate(g)
        m.apply_saved("median")


def test_staleness_policy():
    p = StalenessPolicy(4, "drop")
    assert [p.weight(s) for s in (0, 4, 5, -1)] == [1.0, 1.0, 0.0, 0.0]
    q = StalenessPolicy(2, "scale")
    assert q.weight(1) == 0.5 and q.weight(3) == 0.0


def test_client_helpers():
    assert concat_with_empty(None, torch.ones(2, 3)).shape == (2, 3)
    assert concat_with_empty(torch.zeros(0, 3), torch.ones(2, 3)).shape == (2, 3)
    assert slice_with_empty(torch.arange(10), 8, 5).tolist() == [8, 9]
    buf = add_rows(None, torch.ones(3), [3])                                            # one example
    buf = add_rows(buf, torch.zeros(2, 3), [3])                                         # a batch
    assert buf.shape == (3, 3)
    with pytest.raises(ValueError):
        add_rows(buf, torch.zeros(2, 4), [3])
    set_cookie("k", "v")
    assert get_cookie("k") == "v" and get_cookie("missing") is None


def test_client_id_is_kept_between_clients_and_config_is_strict():
    hub = LoopbackHub()
    from mock_model import MockModel
    sm = MockModel([torch.zeros(1)])
    sm.version = "v"
    FederatedServer(hub, sm, {"serverHyperparams": {"minUpdatesPerVersion": 1}}).setup()
    a = FederatedClient(hub.client, MockModel([torch.zeros(1)]))
    b = FederatedClient(hub.client, MockModel([torch.zeros(1)]))
    assert a.clientId == b.clientId                                                     # cookie-like reuse
    c = FederatedClient(hub.client, MockModel([torch.zeros(1)]), {"clientId": "mine"})
    assert c.clientId == "mine"
    with pytest.raises(ValueError):
        FederatedClient(hub.client, MockModel([torch.zeros(1)]), {"nope": 1})
    with pytest.raises(ValueError):
        FederatedServer(hub, sm, {"nope": 1})


def test_connect_and_upload_timeouts():
    with pytest.raises(ConnectTimeout):
        TcpClientEndpoint("tcp://127.0.0.1:9").connect(timeout=0.2)
    hub = LoopbackHub()
    c = hub.client()
    c.connect()
    hub.pump()
    with pytest.raises(AckTimeout):                                                     # nobody registered an Upload handler
        c.emit("uploadVars", {}, 0.1)
    with pytest.raises(ConnectTimeout):
        c.wait_event("downloadVars", 0.1)


def test_batch_norm_nhwc_cpu_matches_torch_and_fuses_add_relu():
    """`ops.batch_norm_nhwc` (CPU statement of the fused CUDA op): BN over the last axis + residual + ReLU, running stats."""
    import torch.nn.functional as F
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x, r = torch.randn(4, 5, 5, 8), torch.randn(4, 5, 5, 8)
    g, b = torch.rand(8) + 0.5, torch.randn(8)
    rm, rv = torch.zeros(8), torch.ones(8)
    y = ops.batch_norm_nhwc(x, g, b, rm, rv, True, 0.1, 1e-5, "relu", r)
    rm2, rv2 = torch.zeros(8), torch.ones(8)
    ref = F.relu(F.batch_norm(x.permute(0, 3, 1, 2), rm2, rv2, g, b, True, 0.1, 1e-5).permute(0, 2, 3, 1) + r)
    assert torch.allclose(y, ref, atol=1e-6) and torch.allclose(rm, rm2) and torch.allclose(rv, rv2)
    ye = ops.batch_norm_nhwc(x, g, b, rm, rv, False, 0.1, 1e-5)
    assert torch.allclose(ye, F.batch_norm(x.permute(0, 3, 1, 2), rm2, rv2, g, b, False, 0.1, 1e-5).permute(0, 2, 3, 1), atol=1e-6)


def test_resnet20_cpu_forward_backward_and_keras_momentum():
    from distriflow_b200.models.resnet import resnet20, resnet_loss
    torch.manual_seed(0)
    m = resnet20()
    x, y = torch.randn(2, 32, 32, 3), torch.randint(0, 10, (2,))
    loss = resnet_loss(m, x, y)
    loss.backward()
    assert torch.isfinite(loss) and all(p.grad is not None for p in m.parameters())
    bn = m.stem.bn                       # Keras momentum 0.9 keeps 90 % of the old moving statistic
    assert bn.momentum == 0.9 and not torch.equal(bn.moving_mean, torch.zeros_like(bn.moving_mean))
    m.eval()
    with torch.no_grad():
        assert m(x).shape == (2, 10)


def test_converter_style_checkpoint_with_manifest_groups_and_shards(tmp_path):
    """The layout of the reference's `experiment/mnist/model.json`: topology wrapped with
    `keras_version` / `backend` / `model_config`, several manifest groups, each with its own shard files."""
    import json
    import numpy as np
    from distriflow_b200 import load_layers_model
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.models.topology import topology_from_model
    torch.manual_seed(0)
    src = mnist_mlp()
    topo = topology_from_model(src)
    named = [(n, p.detach().numpy()) for n, p in
             [(f"{layer.layer_name}/{k}", getattr(layer, k)) for layer in src.layers if hasattr(layer, "kernel")
              for k in ("kernel", "bias")]]
    groups = [named[:2], named[2:]]
    manifest = []
    for gi, grp in enumerate(groups, 1):
        blob = b"".join(np.ascontiguousarray(a, dtype="<f4").tobytes() for _, a in grp)
        half = len(blob) // 2 // 4 * 4
        paths = [f"group{gi}-shard1of2", f"group{gi}-shard2of2"]          # a tensor may straddle two shards
        (tmp_path / paths[0]).write_bytes(blob[:half])
        (tmp_path / paths[1]).write_bytes(blob[half:])
        manifest.append({"paths": paths, "weights": [{"name": n, "shape": list(a.shape), "dtype": "float32"} for n, a in grp]})
    wrapped = {"keras_version": "2.1.4", "backend": "tensorflow", "model_config": topo.get("model_config", topo),
               "training_config": {"loss": "categorical_crossentropy"}}
    (tmp_path / "model.json").write_text(json.dumps({"modelTopology": wrapped, "weightsManifest": manifest}))
    dst = load_layers_model(str(tmp_path))
    x = torch.rand(3, 28, 28, 1)
    assert torch.allclose(dst(x), src(x), atol=1e-6)
    # a manifest that promises more bytes than the shards hold is an error, not a silent truncation
    manifest[1]["weights"][-1]["shape"] = [1000]
    (tmp_path / "model.json").write_text(json.dumps({"modelTopology": wrapped, "weightsManifest": manifest}))
    with pytest.raises(ValueError):
        load_layers_model(str(tmp_path))
