"""Real multi-GPU runs (2+ B200s on one box): fused reduce-scatter+optimizer+all-gather vs NCCL all-reduce + PyTorch
optimizer, async invariants with concurrent ranks, work queue shared by ranks."""
import os
import socket

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sync_worker(rank, world, port, out, variants=("p2p", "nccl", "p2p-graph")):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    cfg = GPT2Config.tiny()
    res = {}
    for backend in variants:
        torch.manual_seed(0)
        model = GPT2(cfg).to(dev)
        eng = DataParallelEngine(model, gpt2_loss, {"optimizer": "adam", "learningRate": 1e-3, "bucketBytes": 1 << 18,
                                                    "backend": backend.split("-")[0], "cudaGraph": "graph" in backend.split("-"),
                                                    "nvls": "nvls" in backend.split("-")},
                                 rank=rank, world=world, device=dev)
        g = torch.Generator().manual_seed(100 + rank)            # different data per rank
        losses = []
        idx = torch.randint(0, cfg.vocab_size, (4, 65), generator=g).to(dev)     # one fixed batch per rank: loss must fall
        for _ in range(6):
            losses.append(float(eng.step(idx[:, :-1].contiguous(), idx[:, 1:].contiguous())))
        torch.cuda.synchronize()
        if backend.startswith("p2p"):
            with torch.no_grad():                      # one more forward: the fused all-gather GEMM refreshes its kernel
                gpt2_loss(model, idx[:, :-1].contiguous(), idx[:, 1:].contiguous())
            torch.cuda.synchronize()
            masters = eng.full_precision_weights()
            live = [p for p in model.parameters() if p.requires_grad]
            res["working_equals_master"] = all(torch.equal(p.data, m.to(p.dtype)) for p, m in zip(live, masters))
            res["fused"] = getattr(model.first_dense_kernel(), "_dfb_ag", None) is not None
        w = torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])
        res[backend] = (losses, w.cpu())
        if "w_init" not in res:
            torch.manual_seed(0)
            res["w_init"] = torch.cat([p.detach().float().reshape(-1) for p in GPT2(cfg).parameters() if p.requires_grad]).cpu()
        ws = [torch.empty_like(w) for _ in range(world)]
        dist.all_gather(ws, w)
        res[backend + "_same"] = all(torch.equal(ws[0], t) for t in ws)
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sync_engine_p2p_matches_nccl(world, tmp_path):
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_sync_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = torch.load(out)
    assert res["p2p_same"] and res["nccl_same"], "replicas diverged"
    lp, wp = res["p2p"]
    ln, wn = res["nccl"]
    assert lp[-1] < lp[0]
    assert max(abs(a - b) for a, b in zip(lp, ln)) < 5e-2, (lp, ln)
    # bf16 gradients, fp32 pull-reduce vs NCCL's bf16 ring: the two trajectories may differ by a fraction of the distance the
    # weights actually moved, no more (an engine that applied no update at all fails this: its distance is the whole movement)
    moved = float((wn - res["w_init"]).abs().max())
    assert moved > 1e-3, moved
    assert float((wp - wn).abs().max()) <= 0.2 * moved, (float((wp - wn).abs().max()), moved)
    assert res["fused"] and res["working_equals_master"], "a rank's working weights differ from the gathered masters"
    lg, wg = res["p2p-graph"]                                # same engine replayed from a CUDA graph (cross-rank barriers inside)
    assert res["p2p-graph_same"] and max(abs(a - b) for a, b in zip(lp, lg)) < 2e-2 and float((wp - wg).abs().max()) < 1e-2


def test_sync_engine_nvls_matches_p2p(tmp_path):
    """Gradients reduced inside the NVSwitch (`multimem.ld_reduce`) and weights broadcast by it (`multimem.st`): same training
    trajectory as the P2P pull/push kernel up to the bf16 rounding of the reduced gradient; eager and CUDA-graph replay."""
    import torch.multiprocessing as mp
    from distriflow_b200.parallel.symm import multicast_supported
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    if not multicast_supported(torch.device("cuda", 0)):
        pytest.skip("device does not support multicast objects")
    out = str(tmp_path / "res.pt")
    mp.spawn(_sync_worker, args=(world, _free_port(), out, ("p2p", "p2p-nvls", "p2p-nvls-graph")), nprocs=world, join=True)
    res = torch.load(out)
    lp, wp = res["p2p"]
    for v in ("p2p-nvls", "p2p-nvls-graph"):
        lv, wv = res[v]
        assert res[v + "_same"], "replicas diverged"
        assert lv[-1] < lv[0]
        assert max(abs(a - b) for a, b in zip(lp, lv)) < 5e-2, (lp, lv)
        assert float((wp - wv).abs().max()) < 2e-2
    assert res["fused"] and res["working_equals_master"]


def _async_worker(rank, world, port, out, graph=False):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_cnn
    from distriflow_b200.parallel.async_engine import AsyncEngine
    torch.manual_seed(0)
    net = mnist_cnn(10, softmax_output=False, dropout=False).to(dev)
    ce = get_loss("softmaxCrossEntropy")
    eng = AsyncEngine(net, lambda m, x, y: ce(y, m(x)), {"optimizer": "sgd", "learningRate": 1e-3, "maximumStaleness": 4,
                                                          "dtype": "fp32", "gradDtype": "fp32", "cudaGraph": graph},
                      rank=rank, world=world, device=dev)
    g = torch.Generator().manual_seed(rank)
    steps = 12
    losses = []
    for _ in range(steps):
        x = torch.rand(32, 28, 28, 1, generator=g).to(dev)
        y = torch.randint(0, 10, (32,), generator=g).to(dev)
        losses.append(float(eng.step(x, y)))
    eng.drain()
    c = eng.counters()
    allc = [None] * world
    dist.all_gather_object(allc, c)
    if rank == 0:
        torch.save({"counters": allc, "steps": steps, "losses": losses}, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("graph", [False, True])
def test_async_engine_invariants_two_gpus(tmp_path, graph):
    import torch.multiprocessing as mp
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_async_worker, args=(world, _free_port(), out, graph), nprocs=world, join=True)
    res = torch.load(out)
    for c in res["counters"]:
        # every push to this shard was either applied or refused by the staleness guard: nothing lost
        assert c["applied"] + c["dropped"] == world * res["steps"], c
        assert c["version"] == c["applied"] and c["inflight"] == [0, 0] and c["arrived"] == [0, 0]
        assert c["max_staleness_seen"] <= 4


def _stress_worker(rank, world, port, out):
    """Cross-GPU signalling stress: many small buckets back to back, random per-rank delays before each launch, a
    sequence-numbered audit of every result."""
    import random
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from distriflow_b200.ops import _lib
    from distriflow_b200.ops._lib import check, ptr
    from distriflow_b200.parallel.engine import _comm_sigs
    from distriflow_b200.parallel.symm import SymmetricHeap
    L = _comm_sigs()
    n = 64 * 64 * world
    shard = n // world
    heap = SymmetricHeap(n * 8 + 65536, rank, world, None, dev)
    go, po = heap.alloc(n * 4), heap.alloc(n * 4)
    g = heap.view(go, (n,), torch.float32)
    p = heap.view(po, (n,), torch.float32)
    master = torch.zeros(shard, device=dev)
    rng = random.Random(1234 + rank)
    bad = 0
    iters = 300
    for it in range(1, iters + 1):
        g.fill_(float(it * (rank + 1)))                      # gradient of rank r at iteration it = it * (r + 1)
        if rng.random() < 0.5:
            torch.cuda._sleep(rng.randint(1000, 400000))     # desynchronise the ranks on the device
        check(L.dfb_sync_rs_opt_ag(heap.ptrs(go), heap.ptrs(po), heap.ptrs(0), ptr(master), None, None, None, n, shard, rank, world,
                                   it, 1, 0, 0, 0, 1.0, 0.0, 0.9, 0.999, 1e-8, 0.0, it, 8, 0, 0, _lib.stream()), "sync")
        # after the kernel every rank's full weight buffer must hold -sum_{j<=it} j * mean_r(r+1)
        want = -(it * (it + 1) / 2) * (sum(range(1, world + 1)) / world)
        bad += int((p != want).sum().item())
    torch.cuda.synchronize()
    res = torch.tensor([bad], device=dev)
    dist.all_reduce(res)
    if rank == 0:
        torch.save({"bad": int(res), "iters": iters}, out)
    dist.destroy_process_group()


def test_sync_kernel_signalling_stress(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_stress_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = torch.load(out)
    assert res["bad"] == 0, res
