"""Models and engines end to end on one B200: numerics against the CPU oracle / a PyTorch restatement."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda", 0)


def test_mnist_mlp_fit_matches_oracle_on_gpu():
    """K1/K3/K4/K5 on the reference's own shapes ([32,784]x[784,10], N=10) through the hand-written kernels."""
    from distriflow_b200 import DistriTorchModel
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.oracle import MlpOracle, synthetic_mnist
    x, y = synthetic_mnist(32, seed=1)
    for loss in ("referenceCompat", "referenceCompatTrueGrad", "softmaxCrossEntropy"):
        oracle = MlpOracle(seed=3, loss=loss, lr=0.001)
        net = mnist_mlp(softmax_output=(loss != "softmaxCrossEntropy")).to(dev())
        with torch.no_grad():
            for p, v in zip(net.parameters(), oracle.vars):
                p.copy_(torch.from_numpy(v))
        m = DistriTorchModel(net, {"loss": loss, "learningRate": 0.001})
        grads = m.fit(torch.from_numpy(x).to(dev()), torch.from_numpy(y).to(dev()))
        ref_loss, ref_grads = oracle.fit(x, y)
        assert abs(float(m.lastLoss) - ref_loss) < 1e-3 * max(1.0, abs(ref_loss)), loss
        for g, r in zip(grads, ref_grads):
            scale = max(float(np.abs(r).max()), 1e-6)
            assert float((g.cpu() - torch.from_numpy(r)).abs().max()) / scale < 2e-3, loss
        m.update(grads)
        oracle.update(ref_grads)
        for p, v in zip(m.get_vars(), oracle.vars):
            assert np.allclose(p.cpu().numpy(), v, rtol=1e-3, atol=1e-5)


def test_mnist_cnn_forward_backward_gpu_vs_cpu():
    from distriflow_b200.models import mnist_cnn
    torch.manual_seed(0)
    cpu = mnist_cnn(5, dropout=False)
    gpu = mnist_cnn(5, dropout=False)
    gpu.load_state_dict(cpu.state_dict())
    gpu.to(dev())
    x = torch.rand(4, 28, 28, 1)
    y = F.one_hot(torch.randint(0, 5, (4,)), 5).float()
    lc = (-(y * torch.log(cpu(x).clamp_min(1e-7))).sum(-1)).mean()
    lc.backward()
    lg = (-(y.to(dev()) * torch.log(gpu(x.to(dev())).clamp_min(1e-7))).sum(-1)).mean()
    lg.backward()
    assert abs(float(lc) - float(lg)) < 1e-4
    for a, b in zip(cpu.parameters(), gpu.parameters()):
        assert float((a.grad - b.grad.cpu()).abs().max()) < 1e-3 * max(1.0, float(a.grad.abs().max()))


def _torch_gpt2_reference(model, idx, tgt):
    """Plain PyTorch fp32 restatement of models/gpt2.py on the same weights."""
    c = model.config
    p = {n: t.float() for n, t in model.named_parameters()}
    B, T = idx.shape
    x = p["wte"][idx] + p["wpe"][:T]
    for i in range(c.n_layer):
        pre = f"h.{i}."
        h = F.layer_norm(x, (c.n_embd,), p[pre + "ln_1.gamma"], p[pre + "ln_1.beta"], 1e-5)
        qkv = h @ p[pre + "c_attn.kernel"] + p[pre + "c_attn.bias"]
        q, k, v = qkv.view(B, T, 3, c.n_head, c.n_embd // c.n_head).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, c.n_embd)
        x = x + a @ p[pre + "c_proj.kernel"] + p[pre + "c_proj.bias"]
        h = F.layer_norm(x, (c.n_embd,), p[pre + "ln_2.gamma"], p[pre + "ln_2.beta"], 1e-5)
        h = F.gelu(h @ p[pre + "c_fc.kernel"] + p[pre + "c_fc.bias"], approximate="tanh")
        x = x + h @ p[pre + "c_proj2.kernel"] + p[pre + "c_proj2.bias"]
    x = F.layer_norm(x, (c.n_embd,), p["ln_f.gamma"], p["ln_f.beta"], 1e-5)
    return F.cross_entropy((x @ p["wte"].t()).reshape(B * T, -1), tgt.reshape(-1))


def test_gpt2_tiny_loss_and_grads_vs_torch():
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    torch.manual_seed(0)
    cfg = GPT2Config.tiny()
    model = GPT2(cfg).to(dev()).to(torch.bfloat16)
    idx = torch.randint(0, cfg.vocab_size, (4, 64), device=dev())
    tgt = torch.randint(0, cfg.vocab_size, (4, 64), device=dev())
    loss = gpt2_loss(model, idx, tgt)
    loss.backward()
    ref_params = {n: t.detach().float().requires_grad_(True) for n, t in model.named_parameters()}

    class Shim:
        config = cfg

        def named_parameters(self):
            return ref_params.items()

    ref = _torch_gpt2_reference(Shim(), idx, tgt)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 3e-2, (float(loss), float(ref))
    worst = 0.0
    for n, p in model.named_parameters():
        g, r = p.grad.float(), ref_params[n].grad
        worst = max(worst, float((g - r).abs().max() / r.abs().max().clamp_min(1e-6)))
    assert worst < 8e-2, worst


@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_engine_single_gpu_matches_torch_engine(opt):
    """world = 1: the fused bucket kernel is the optimizer; compare with the PyTorch restatement ("nccl" back end)."""
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev())
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    losses = {}
    for backend in ("p2p", "nccl"):
        torch.manual_seed(0)
        model = GPT2(cfg).to(dev())
        eng = DataParallelEngine(model, gpt2_loss, {"optimizer": opt, "learningRate": 1e-3 if opt == "adam" else 0.05,
                                                    "bucketBytes": 1 << 18, "backend": backend}, device=dev())
        losses[backend] = [float(eng.step(x, y)) for _ in range(6)]
    a, b = losses["p2p"], losses["nccl"]
    assert a[-1] < a[0]
    assert max(abs(u - v) for u, v in zip(a, b)) < 5e-2, (a, b)


def test_engine_checkpoint_from_master_weights(tmp_path):
    from distriflow_b200 import load_layers_model
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    net = mnist_mlp(softmax_output=False).to(dev())
    ce = get_loss("softmaxCrossEntropy")
    eng = DataParallelEngine(net, lambda m, x, y: ce(y, m(x)), {"optimizer": "sgd", "learningRate": 1e-3, "dtype": "fp32",
                                                                  "gradDtype": "fp32"}, device=dev())
    x = torch.rand(32, 28, 28, 1, device=dev())
    y = torch.randint(0, 10, (32,), device=dev())
    l0 = float(eng.step(x, y))
    for _ in range(20):
        l1 = float(eng.step(x, y))
    assert l1 < l0
    path = eng.save_checkpoint(str(tmp_path / "v1"))
    back = load_layers_model(path)
    for a, b in zip(eng.full_precision_weights(), back.parameters()):
        assert torch.equal(a.cpu(), b.detach())


def test_graft_smoke():
    import __graft_entry__ as g
    g.smoke()


def test_async_engine_with_device_work_queue_single_gpu():
    """world = 1: ticket -> gather -> pull -> fit -> push -> apply -> complete; every batch exactly once per epoch."""
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.oracle import synthetic_mnist
    from distriflow_b200.parallel.async_engine import AsyncEngine
    from distriflow_b200.parallel.workqueue import DeviceWorkQueue
    torch.manual_seed(0)
    x, y = synthetic_mnist(256, seed=2)
    xd, yd = torch.from_numpy(x).to(dev()), torch.from_numpy(y).to(dev())
    net = mnist_mlp(softmax_output=False).to(dev())
    ce = get_loss("softmaxCrossEntropy")
    eng = AsyncEngine(net, lambda m, a, b: ce(b, m(a)), {"optimizer": "sgd", "learningRate": 1e-4, "dtype": "fp32", "gradDtype": "fp32",
                                                         "maximumStaleness": 4}, device=dev())
    wq = DeviceWorkQueue(eng.state.heaps[0], 8, 2)
    losses = []
    n = eng.run_dataset(wq, xd, yd, 32, on_step=lambda i, l: losses.append(float(l)))
    eng.drain()
    c = eng.counters()
    assert c["applied"] == 16 and c["dropped"] == 0 and c["version"] == 16, c      # 8 batches x 2 epochs, nothing lost
    assert 16 <= n <= 16 + 6                                                         # a few no-op steps after the queue drained
    assert losses[15] < losses[0]
    assert wq.take().done


def test_metric_ring():
    from distriflow_b200.utils.metrics import DeviceMetricRing
    ring = DeviceMetricRing(slots=4, device=dev())
    pred = torch.randn(64, 10, device=dev())
    lab = torch.randint(0, 10, (64,), device=dev())
    for i in range(6):
        ring.record(pred, lab, loss=torch.tensor(float(i), device=dev()))
    got = ring.poll(block=True)
    assert len(got) == 6 and got[-1]["loss"] == 5.0
    assert abs(got[0]["accuracy"] - float((pred.argmax(-1) == lab).float().mean())) < 1e-6


@pytest.mark.parametrize("opt", ["momentum", "adam"])
def test_engine_cuda_graph_matches_eager(opt):
    """Whole-step CUDA graph (forward, backward, fused bucket kernels): same losses as the eager engine."""
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    g = torch.Generator().manual_seed(5)
    batches = []
    for _ in range(8):
        idx = torch.randint(0, cfg.vocab_size, (4, 65), generator=g).to(dev())
        batches.append((idx[:, :-1].contiguous(), idx[:, 1:].contiguous()))
    out = {}
    for graph in (False, True):
        torch.manual_seed(0)
        model = GPT2(cfg).to(dev())
        eng = DataParallelEngine(model, gpt2_loss, {"optimizer": opt, "learningRate": 1e-3 if opt == "adam" else 0.05,
                                                    "bucketBytes": 1 << 18, "cudaGraph": graph}, device=dev())
        out[graph] = [float(eng.step(x, y)) for x, y in batches]
        assert eng.version == 8
    assert max(abs(a - b) for a, b in zip(out[False], out[True])) < 2e-2, out


@pytest.mark.parametrize("graph", [False, True])
def test_engine_state_resume_on_gpu(tmp_path, graph):
    """bf16 working weights + fp32 master shards + Adam moments saved and restored: the resumed run continues the original
    one step for step (eager and CUDA-graph mode, where the step counter lives in device memory)."""
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    g = torch.Generator(device="cuda").manual_seed(5)
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev(), generator=g)
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    # `deterministic`: the resumed run must reproduce the original one, so the embedding gradient takes the sorted reduction
    conf = {"optimizer": "adam", "learningRate": 1e-3, "bucketBytes": 1 << 18, "cudaGraph": graph, "deterministic": True}
    torch.manual_seed(0)
    eng = DataParallelEngine(GPT2(cfg).to(dev()), gpt2_loss, conf, device=dev())
    for _ in range(5):
        eng.step(x, y)
    eng.save_state(str(tmp_path))
    cont = [float(eng.step(x, y)) for _ in range(4)]
    torch.manual_seed(123)                                   # different initial weights: the state must replace them
    eng2 = DataParallelEngine(GPT2(cfg).to(dev()), gpt2_loss, conf, device=dev())
    eng2.load_state(str(tmp_path))
    assert eng2.version == 5
    resumed = [float(eng2.step(x, y)) for _ in range(4)]
    assert max(abs(a - b) for a, b in zip(cont, resumed)) < 2e-3, (cont, resumed)
    for a, b in zip(eng.full_precision_weights(), eng2.full_precision_weights()):
        assert float((a - b).abs().max()) < 1e-4


def test_engine_trace_phases_and_watchdog_counter_on_gpu():
    """`trace: True` gives device-timed forward / backward / collective phases; the watchdog's progress word is written by
    the GPU (pinned host memory) at the end of every step."""
    import time
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev())
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    torch.manual_seed(0)
    eng = DataParallelEngine(GPT2(cfg).to(dev()), gpt2_loss, {"optimizer": "adam", "learningRate": 1e-3, "trace": True,
                                                             "watchdogSeconds": 0.3, "cudaGraph": True}, device=dev())
    for _ in range(5):
        eng.step(x, y)
    s = eng.trace_summary()
    assert set(s) == {"forward", "backward", "collective_tail"}
    assert s["forward"]["count"] == 5 and s["forward"]["mean_ms"] > 0 and s["backward"]["mean_ms"] > 0
    torch.cuda.synchronize()
    assert int(eng._prog_pin[0]) == 5
    time.sleep(0.9)
    eng.close()
    assert eng.stall_reports and eng.stall_reports[0][0] == 5


def test_engine_watchdog_counter_inside_cuda_graph():
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev())
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    torch.manual_seed(0)
    eng = DataParallelEngine(GPT2(cfg).to(dev()), gpt2_loss, {"optimizer": "sgd", "learningRate": 1e-2, "watchdogSeconds": 30.0,
                                                             "cudaGraph": True}, device=dev())
    for _ in range(7):                       # 3 eager + capture/replay + 3 replays
        eng.step(x, y)
    torch.cuda.synchronize()
    eng.close()
    assert eng._graph is not None and int(eng._prog_pin[0]) == 7 and eng.stall_reports == []


def test_learning_rate_schedule_reaches_replayed_graph():
    """`set_learning_rate` under `cudaGraph`: the captured bucket kernels read the rate from device memory."""
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev())
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    torch.manual_seed(0)
    eng = DataParallelEngine(GPT2(cfg).to(dev()), gpt2_loss, {"optimizer": "sgd", "learningRate": 0.05, "cudaGraph": True},
                             device=dev())
    for _ in range(5):
        eng.step(x, y)
    assert eng._graph is not None
    w0 = torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])
    eng.set_learning_rate(0.0)
    for _ in range(2):
        eng.step(x, y)
    w1 = torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])
    assert torch.equal(w0, w1), "rate 0 must freeze the weights in replayed steps"
    eng.set_learning_rate(0.05)
    eng.step(x, y)
    w2 = torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])
    assert not torch.equal(w1, w2)


def test_trainer_on_gpu_with_host_data_prefetch_and_resume(tmp_path):
    """`Trainer` over host-resident data: batches reach the GPU through `DevicePrefetcher` (pinned memory, copy stream); an
    interrupted run resumed from `<saveDir>/current` ends with the weights of the uninterrupted one."""
    from distriflow_b200 import Trainer
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.parallel.engine import DataParallelEngine
    from distriflow_b200.utils.prefetch import DevicePrefetcher
    g = torch.Generator().manual_seed(1)
    x = torch.rand(96, 28, 28, 1, generator=g)
    y = torch.randint(0, 10, (96,), generator=g)
    pf = DevicePrefetcher([(x[i:i + 32], y[i:i + 32]) for i in range(0, 96, 32)], dev(), depth=2)
    got = list(pf)
    assert len(got) == 3 and all(t.device == dev() for b in got for t in b) and pf.h2d_bytes == x.numel() * 4 + y.numel() * 8
    torch.cuda.synchronize()
    assert torch.equal(torch.cat([b[0] for b in got]).cpu(), x) and torch.equal(torch.cat([b[1] for b in got]).cpu(), y)

    ce = get_loss("softmaxCrossEntropy")
    conf = {"optimizer": "momentum", "learningRate": 0.01, "dtype": "fp32", "gradDtype": "fp32"}
    tconf = {"batchSize": 32, "epochs": 4, "checkpointEvery": 5}

    def make(seed):
        torch.manual_seed(seed)
        return DataParallelEngine(mnist_mlp(softmax_output=False).to(dev()), lambda m, a, b: ce(b, m(a)), conf, device=dev())

    ref = make(0)
    Trainer(ref, x, y, tconf).fit()
    first = make(0)
    out1 = Trainer(first, x, y, {**tconf, "saveDir": str(tmp_path)}).fit(max_steps=7)
    assert out1["steps"] == 7 and not out1["finished"]
    second = make(9)
    out2 = Trainer(second, x, y, {**tconf, "saveDir": str(tmp_path)}).fit()
    assert out2["resumed_from"] == 7 and out2["steps"] == 12 and out2["finished"]
    for a, b in zip(ref.full_precision_weights(), second.full_precision_weights()):
        assert float((a - b).abs().max()) < 1e-5


@pytest.mark.parametrize("model_name", ["gpt2", "mnist_cnn", "resnet20"])
def test_no_weight_is_read_after_its_bucket_is_reported(model_name):
    """Every parameter in its own bucket + `debugPoison`: the engine overwrites a bucket's working weights with NaN the
    moment its last gradient is reported (which is when the overlapped fused kernel may start rewriting them).  A backward
    function that reports a weight and reads it afterwards (the round-1 `_MlpFn` / `_ThinConvFn` bug) yields NaN gradients,
    which this test sees as a non-finite or diverging loss against the plain run."""
    from distriflow_b200.parallel.engine import DataParallelEngine
    if model_name == "gpt2":
        from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
        cfg = GPT2Config.tiny()
        idx = torch.randint(0, cfg.vocab_size, (4, 65), device=dev())
        batch = (idx[:, :-1].contiguous(), idx[:, 1:].contiguous())
        make, loss_fn = (lambda: GPT2(cfg).to(dev())), gpt2_loss
    elif model_name == "mnist_cnn":
        from distriflow_b200.losses import get_loss
        from distriflow_b200.models import mnist_cnn
        ce = get_loss("softmaxCrossEntropy")
        loss_fn = lambda m, a, b: ce(b, m(a))                                     # noqa: E731
        make = lambda: mnist_cnn(5, dropout=False, softmax_output=False).to(dev())      # noqa: E731
        batch = (torch.rand(32, 28, 28, 1, device=dev()), torch.randint(0, 5, (32,), device=dev()))
    else:
        from distriflow_b200.models.resnet import resnet20, resnet_loss as loss_fn
        make = lambda: resnet20().to(dev())                                       # noqa: E731
        batch = (torch.rand(64, 32, 32, 3, device=dev()), torch.randint(0, 10, (64,), device=dev()))
    dt = "fp32" if model_name == "mnist_cnn" else "bf16"
    out = {}
    for poison in (False, True):
        torch.manual_seed(0)
        eng = DataParallelEngine(make(), loss_fn, {"optimizer": "sgd", "learningRate": 1e-2, "bucketBytes": 1, "dtype": dt,
                                                   "gradDtype": dt, "debugPoison": poison, "fuseAllGather": False}, device=dev())
        assert len(eng.buckets) == sum(1 for _ in eng.model.parameters())
        out[poison] = [float(eng.step(*batch)) for _ in range(4)]
    assert all(np.isfinite(v) for v in out[True]), out
    assert max(abs(a - b) for a, b in zip(out[False], out[True])) < 2e-2 * max(1.0, abs(out[False][0])), out
