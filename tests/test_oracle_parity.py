"""Non-degenerate parity with the CPU oracle: WEIGHTS, not only the loss, after a run of
the reference's schedule through `DistriServer(backend="b200")`, for the raw 0..255 inputs of the reference
(`/root/reference/experiment/mnist/mnist_data.ts:41-44`, where the model's softmax saturates and the loss takes few values)
and for inputs / 255 where the loss really falls; MLP (`mnist_server.ts:16-22`) and CNN (`experiment/mnist/model.json`)."""
import numpy as np
import pytest
import torch

from distriflow_b200 import DistriDataset, DistriServer
from distriflow_b200.losses import get_loss
from distriflow_b200.models import mnist_cnn, mnist_mlp
from distriflow_b200.oracle import CnnOracle, MlpOracle, simulate_async, synthetic_mnist

DEVICES = [pytest.param(torch.device("cpu"), id="cpu"),
           pytest.param(torch.device("cuda", 0), id="b200", marks=pytest.mark.gpu)]


@pytest.mark.parametrize("loss", ["softmaxCrossEntropy", "referenceCompat", "referenceCompatTrueGrad"])
def test_cnn_oracle_gradients_match_autograd(loss):
    x, y = synthetic_mnist(6, classes=5, seed=3)
    x = x / 255.0
    o = CnnOracle(classes=5, seed=1, loss=loss)
    net = mnist_cnn(5, softmax_output=(loss != "softmaxCrossEntropy"), dropout=False)
    with torch.no_grad():
        for p, v in zip(net.parameters(), o.vars):
            p.copy_(torch.from_numpy(v))
    out = get_loss(loss)(torch.from_numpy(y), net(torch.from_numpy(x).float()))
    out.backward()
    ref_loss, grads = o.fit(x, y)
    assert abs(float(out) - ref_loss) < 1e-5
    for p, g in zip(net.parameters(), grads):
        assert float((p.grad - torch.from_numpy(g)).abs().max()) <= 1e-4 * max(float(np.abs(g).max()), 1e-6)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("model", ["mlp", "cnn"])
@pytest.mark.parametrize("normalise", [False, True], ids=["raw255", "div255"])
def test_weights_after_training_match_the_oracle(device, model, normalise):
    if model == "cnn" and not normalise:
        # raw 0..255 pixels through two conv layers give activations of 1e5 and a saturated softmax: the update is noise that
        # depends on fp32 summation order (the reference runs only its dense model on raw pixels, mnist_server.ts:14,26)
        pytest.skip("ill-conditioned: no oracle-grade comparison exists for the CNN on un-normalised pixels")
    steps, bs, lr = (48, 32, 0.001) if model == "mlp" else (8, 16, 0.001)
    x, y = synthetic_mnist(steps * bs, seed=0)
    if normalise:
        x = x / 255.0
        lr = 0.05 if model == "mlp" else 0.02
    oracle = (MlpOracle(seed=1, loss="referenceCompat", lr=lr) if model == "mlp"
              else CnnOracle(classes=10, seed=1, loss="referenceCompat", lr=lr))
    net = mnist_mlp(softmax_output=True) if model == "mlp" else mnist_cnn(10, softmax_output=True, dropout=False)
    with torch.no_grad():
        for p, v in zip(net.parameters(), oracle.vars):
            p.copy_(torch.from_numpy(v))
    ds = DistriDataset(torch.from_numpy(x), torch.from_numpy(y), {"batchSize": bs, "epochs": 1})
    server = DistriServer(None, net, ds, {"modelCompileConfig": {"loss": "referenceCompat", "learningRate": lr},
                                          "serverHyperparams": {"minUpdatesPerVersion": 1}, "sendMetrics": True,
                                          "engine": {"dtype": "fp32", "gradDtype": "fp32"}},
                          mode="sync", backend="b200", device=device)
    losses = []
    server.on_upload(lambda m: losses.append(m.metrics[0]))
    server.run()
    ref = simulate_async(oracle, x, y, batch_size=bs, epochs=1, workers=1)
    assert len(losses) == len(ref["loss"]) == steps
    tol = 1e-6 if device.type == "cpu" else 2e-5
    moved = max(float(np.abs(v - v0).max()) for v, v0 in zip(oracle.vars, ref["initial_vars"]))
    assert moved > (1e-4 if normalise else 0.0)
    for w, v in zip(server.model.get_vars(), oracle.vars):
        assert np.allclose(w.cpu().numpy(), v, rtol=0, atol=max(tol, 1e-3 * moved)), float(np.abs(w.cpu().numpy() - v).max())
    assert max(abs(a - b) for a, b in zip(losses, ref["loss"])) < 1e-4
    if normalise and model == "mlp":
        assert ref["loss"][-1] < ref["loss"][0] - 1e-3            # un-saturated softmax: the curve really falls
