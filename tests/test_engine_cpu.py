"""Synchronous data-parallel engine on CPU: `gloo` plumbing with world_size 1 and 2, optimizer parity with torch.optim,
resume from a saved state (also across world sizes).

Reference behaviour being replaced: FederatedServer's "collect minUpdatesPerVersion gradients, mean, update, broadcast"
(`/root/reference/src/server/federated_server.ts:71-116`) and the version-directory checkpoint
(`/root/reference/src/server/models.ts:132-146`).  The CUDA path of the same engine is covered by tests/test_gpu_engine.py
and tests/test_multigpu.py.
"""
import os
import socket

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F


def _net(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(12, 16), nn.Tanh(), nn.Linear(16, 4))


def _loss(m, x, y):
    return F.cross_entropy(m(x), y)


def _batch(seed, n=8):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 12, generator=g), torch.randint(0, 4, (n,), generator=g)


@pytest.mark.parametrize("opt", ["sgd", "momentum", "adam"])
def test_engine_cpu_matches_torch_optim(opt):
    from distriflow_b200.parallel.engine import DataParallelEngine
    net, ref = _net(1), _net(1)
    eng = DataParallelEngine(net, _loss, {"optimizer": opt, "learningRate": 0.05, "momentum": 0.9, "dtype": "fp32",
                                          "gradDtype": "fp32"}, device=torch.device("cpu"))
    topt = {"sgd": lambda p: torch.optim.SGD(p, lr=0.05), "momentum": lambda p: torch.optim.SGD(p, lr=0.05, momentum=0.9),
            "adam": lambda p: torch.optim.Adam(p, lr=0.05)}[opt](ref.parameters())
    for s in range(5):
        x, y = _batch(s)
        l_e = float(eng.step(x, y))
        topt.zero_grad()
        l_r = _loss(ref, x, y)
        l_r.backward()
        topt.step()
        assert abs(l_e - float(l_r)) < 1e-5
    for a, b in zip(eng.full_precision_weights(), ref.parameters()):
        assert torch.allclose(a, b.detach(), atol=1e-5)
    assert eng.version == 5


def test_engine_state_roundtrip_resumes_exactly(tmp_path):
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = {"optimizer": "adam", "learningRate": 0.01, "dtype": "fp32", "gradDtype": "fp32"}
    eng = DataParallelEngine(_net(2), _loss, cfg, device=torch.device("cpu"))
    for s in range(3):
        eng.step(*_batch(s))
    path = eng.save_state(str(tmp_path))
    assert path and os.path.exists(path) and not os.path.exists(path + ".tmp")
    cont = [float(eng.step(*_batch(10 + s))) for s in range(3)]
    eng2 = DataParallelEngine(_net(99), _loss, cfg, device=torch.device("cpu"))      # different init: must be overwritten
    eng2.load_state(str(tmp_path))
    assert eng2.version == 3
    resumed = [float(eng2.step(*_batch(10 + s))) for s in range(3)]
    assert cont == resumed
    for a, b in zip(eng.full_precision_weights(), eng2.full_precision_weights()):
        assert torch.equal(a, b)
    with pytest.raises(ValueError):
        eng2.load_state_dict({"format": "something else"})
    sd = eng.state_dict()
    sd["params"].pop(next(iter(sd["params"])))
    with pytest.raises(ValueError):
        eng2.load_state_dict(sd)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = {"optimizer": "momentum", "learningRate": 0.05, "dtype": "fp32", "gradDtype": "fp32", "backend": "gloo",
           "bucketBytes": 512}                                         # several buckets
    eng = DataParallelEngine(_net(3 + rank), _loss, cfg, rank=rank, world=world, device=torch.device("cpu"))   # rank 0's init wins
    losses = []
    for s in range(4):
        x, y = _batch(100 + s, 16)
        lo = rank * 8
        losses.append(float(eng.step(x[lo:lo + 8], y[lo:lo + 8])))     # each rank: its half of the global batch
    eng.save_state(out_dir)
    w = torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])
    ws = [torch.empty_like(w) for _ in range(world)]
    dist.all_gather(ws, w)
    if rank == 0:
        torch.save({"w": w, "same": all(torch.equal(ws[0], t) for t in ws), "losses": losses}, os.path.join(out_dir, "res.pt"))
    dist.destroy_process_group()


def test_engine_gloo_two_ranks_equal_one_rank_on_the_global_batch(tmp_path):
    import torch.multiprocessing as mp
    from distriflow_b200.parallel.engine import DataParallelEngine
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    res = torch.load(os.path.join(str(tmp_path), "res.pt"))
    assert res["same"], "replicas diverged"
    # one process on the whole batch: mean of the two half-batch gradients = gradient of the mean loss
    cfg = {"optimizer": "momentum", "learningRate": 0.05, "dtype": "fp32", "gradDtype": "fp32"}
    solo = DataParallelEngine(_net(3), _loss, cfg, device=torch.device("cpu"))
    for s in range(4):
        solo.step(*_batch(100 + s, 16))
    w = torch.cat([t.reshape(-1) for t in solo.full_precision_weights()])
    assert torch.allclose(w, res["w"], atol=1e-5)
    # the state saved by the 2-rank run resumes in a 1-rank engine (world-size independent format)
    again = DataParallelEngine(_net(50), _loss, cfg, device=torch.device("cpu"))
    again.load_state(str(tmp_path))
    assert again.version == 4
    w2 = torch.cat([t.reshape(-1) for t in again.full_precision_weights()])
    assert torch.equal(w2, res["w"])
    a, b = float(again.step(*_batch(200, 16))), float(solo.step(*_batch(200, 16)))
    assert abs(a - b) < 1e-5


def test_set_learning_rate_takes_effect_on_cpu():
    from distriflow_b200.parallel.engine import DataParallelEngine
    eng = DataParallelEngine(_net(4), _loss, {"optimizer": "sgd", "learningRate": 0.1, "dtype": "fp32", "gradDtype": "fp32"},
                             device=torch.device("cpu"))
    eng.step(*_batch(0))
    w0 = [t.clone() for t in eng.full_precision_weights()]
    eng.set_learning_rate(0.0)
    eng.step(*_batch(1))
    assert all(torch.equal(a, b) for a, b in zip(w0, eng.full_precision_weights()))
    eng.set_learning_rate(0.1)
    eng.step(*_batch(2))
    assert not all(torch.equal(a, b) for a, b in zip(w0, eng.full_precision_weights()))


def test_tied_weight_reports_once_per_producer_and_matches_plain_autograd():
    """GPT-2's `wte` feeds the embedding gather AND the LM head (`_dfb_grad_parts = 2`).  On the plain-autograd path both
    gradients are accumulated by autograd and delivered by ONE hook call, which must report both parts: the bucket is
    launched, and the step equals a plain PyTorch step on the same weights."""
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    torch.manual_seed(0)
    model = GPT2(cfg)
    assert getattr(model.wte, "_dfb_grad_parts", 1) == 2
    torch.manual_seed(0)
    ref = GPT2(cfg)
    opt = torch.optim.SGD([p for p in ref.parameters() if p.requires_grad], lr=0.1)
    eng = DataParallelEngine(model, gpt2_loss, {"optimizer": "sgd", "learningRate": 0.1, "dtype": "fp32", "gradDtype": "fp32"},
                             device=torch.device("cpu"))
    g = torch.Generator().manual_seed(1)
    idx = torch.randint(0, cfg.vocab_size, (2, 33), generator=g)
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    for _ in range(3):
        l_eng = float(eng.step(x, y))
        assert all(b.pending == 0 for b in eng.buckets)
        opt.zero_grad()
        l_ref = gpt2_loss(ref, x, y)
        l_ref.backward()
        opt.step()
        assert abs(l_eng - float(l_ref)) < 1e-5
    for a, b in zip(eng.full_precision_weights(), [p for p in ref.parameters() if p.requires_grad]):
        assert torch.allclose(a, b.detach(), atol=1e-6)


def test_deterministic_key_switches_the_embedding_gradient_path_per_engine():
    """`deterministic: True` selects the sorted embedding-gradient reduction process-wide; the next engine built without it
    switches back (the flag follows the most recently built engine, like the closing-barrier mode of the bucket kernels)."""
    import torch.nn.functional as F
    from distriflow_b200.ops import embed
    from distriflow_b200.parallel.engine import DataParallelEngine
    net = torch.nn.Linear(4, 3)
    loss = lambda m, x, y: F.mse_loss(m(x), y)               # noqa: E731
    DataParallelEngine(net, loss, {"optimizer": "sgd", "learningRate": 0.1, "dtype": "fp32", "gradDtype": "fp32",
                                   "deterministic": True}, device=torch.device("cpu"))
    assert embed.DETERMINISTIC is True
    DataParallelEngine(torch.nn.Linear(4, 3), loss, {"optimizer": "sgd", "learningRate": 0.1, "dtype": "fp32", "gradDtype": "fp32"},
                       device=torch.device("cpu"))
    assert embed.DETERMINISTIC is embed.DETERMINISTIC_ENV
