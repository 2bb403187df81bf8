"""Real multi-GPU runs (2+ B200s on one box): fused reduce-scatter+optimizer+all-gather vs NCCL all-reduce + PyTorch
optimizer, async invariants with concurrent ranks, work queue shared by ranks."""
import os
import socket

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


WORLD = int(os.environ.get("DFB_TEST_WORLD", "2"))        # DFB_TEST_WORLD=8 runs the same tests on a whole box


# The arms of these tests are compared with each other: keep the one order-dependent kernel of the GPT-2 step (the in-place
# embedding-gradient scatter) out of the comparison; spawned workers inherit the variable.  Set at import, so it also holds
# for every other test of the same pytest process (collection imports this file first): the engine tests there compare runs
# with each other too (resume, graph vs eager).  The scatter path itself is covered by
# test_gpu_kernels.py::test_embedding_backward_scatters_into_a_filled_gradient_slice (which switches it back on), by
# smoke() and by every bench run (separate processes without the variable).
os.environ.setdefault("DFB_EMBEDDING_DETERMINISTIC", "1")


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
        flags = backend.split("-")
        bucket = next((int(f[1:]) for f in flags if f[:1] == "b" and f[1:].isdigit()), 1 << 18)
        eng = DataParallelEngine(model, gpt2_loss, {"optimizer": "adam", "learningRate": 1e-3, "bucketBytes": bucket,
                                                    "backend": flags[0], "cudaGraph": "graph" in flags, "nvls": "nvls" in flags,
                                                    "closeBarrier": "bucket" if "bucketclose" in flags else "step"},
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
            res["working_equals_master"] = res.get("working_equals_master", True) and \
                all(torch.equal(p.data, m.to(p.dtype)) for p, m in zip(live, masters))       # every p2p variant, not the last one
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


@pytest.mark.parametrize("world", [WORLD])
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
    # weights actually moved, no more (an engine that applied no update at all fails this: its distance is the whole movement).
    # L2 norms, not the maximum: under Adam a weight whose gradient is rounding noise moves by +-lr per step in either arm.
    moved = float((wn - res["w_init"]).norm())
    assert moved > 1e-2, moved
    assert float((wp - wn).norm()) <= 0.2 * moved, (float((wp - wn).norm()), moved)
    assert res["fused"] and res["working_equals_master"], "a rank's working weights differ from the gathered masters"
    lg, wg = res["p2p-graph"]                                # same engine replayed from a CUDA graph (cross-rank barriers inside)
    assert res["p2p-graph_same"] and max(abs(a - b) for a, b in zip(lp, lg)) < 2e-2 and float((wp - wg).abs().max()) < 1e-2


def test_sync_engine_bucket_sizes_and_barrier_modes(tmp_path):
    """Bucket boundaries on every kind of parameter (64 KB buckets isolate single kernels, one bucket holds the whole model)
    and both closing-barrier protocols: every variant must train like the NCCL arm and keep the replicas bit-identical."""
    import torch.multiprocessing as mp
    world = WORLD
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    variants = ("nccl", "p2p-b65536", "p2p-b1048576", "p2p-b1073741824", "p2p-b65536-bucketclose", "p2p-b65536-graph")
    mp.spawn(_sync_worker, args=(world, _free_port(), out, variants), nprocs=world, join=True)
    res = torch.load(out)
    ln, wn = res["nccl"]
    moved = float((wn - res["w_init"]).norm())
    dist_to_nccl = {v: round(float((res[v][1] - wn).norm()) / moved, 3) for v in variants[1:]}
    print("distance to the NCCL arm / distance moved:", dist_to_nccl)
    for v in variants[1:]:
        lv, wv = res[v]
        assert res[v + "_same"], f"{v}: replicas diverged"
        assert max(abs(a - b) for a, b in zip(lv, ln)) < 5e-2, (v, lv, ln)
        # bf16 + Adam turns rounding noise into +-lr steps; 6 steps from the same start stay within a fraction of the way moved
        assert float((wv - wn).norm()) <= 0.2 * moved, (v, dist_to_nccl)
    assert res["working_equals_master"]


def _bitwise_worker(rank, world, port, out):
    """fp32 gradients + SGD / momentum / Adam through the fused kernel on REAL peer mappings: every rank's weights must equal,
    bit for bit, the same kernel run by one process (R = 1) on gradients summed in rank order."""
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
    L.dfb_sync_close_per_bucket(1)
    n = 64 * 96 * world
    shard = n // world
    heap = SymmetricHeap(n * 8 + 65536, rank, world, None, dev)
    go, po = heap.alloc(n * 4), heap.alloc(n * 4)
    g, p = heap.view(go, (n,), torch.float32), heap.view(po, (n,), torch.float32)
    solo = SymmetricHeap(n * 8 + 65536, 0, 1, None, dev)             # the single-process reference runs on this rank's GPU
    sgo, spo = solo.alloc(n * 4), solo.alloc(n * 4)
    sg, sp = solo.view(sgo, (n,), torch.float32), solo.view(spo, (n,), torch.float32)
    bad = {}
    epoch = 0
    for opt in (0, 1, 2):
        gen = torch.Generator(device="cpu").manual_seed(7)
        w0 = torch.randn(n, generator=gen)
        p.copy_(w0)
        sp.copy_(w0)
        master, m, v = w0[rank * shard:(rank + 1) * shard].to(dev).clone(), torch.zeros(shard, device=dev), torch.zeros(shard, device=dev)
        smaster, sm, sv = w0.to(dev).clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        for step in range(1, 5):
            grads = [torch.randn(n, generator=gen) for _ in range(world)]         # the same list on every rank
            g.copy_(grads[rank])
            epoch += 1
            torch.cuda.synchronize()
            dist.barrier()
            check(L.dfb_sync_rs_opt_ag(heap.ptrs(go), heap.ptrs(po), heap.ptrs(0), ptr(master), ptr(m), ptr(v), None, n, shard, rank,
                                       world, epoch, 1, 0, 0, opt, 0.01, 0.9, 0.9, 0.999, 1e-8, 0.0, step, 16, 0, 0, _lib.stream()),
                  "sync")
            acc = torch.zeros(n)
            for q in range(world):                                                 # the kernel's order: rank 0 first
                acc = acc + grads[q]
            # R = 1 applies `acc * 1.0`: fold 1/R in by giving it the sum scaled exactly as the R-rank kernel scales it
            sg.copy_(acc * torch.tensor(1.0 / world, dtype=torch.float32))
            check(L.dfb_sync_rs_opt_ag(solo.ptrs(sgo), solo.ptrs(spo), solo.ptrs(0), ptr(smaster), ptr(sm), ptr(sv), None, n, n, 0, 1,
                                       0, 1, 0, 0, opt, 0.01, 0.9, 0.9, 0.999, 1e-8, 0.0, step, 16, 0, 0, _lib.stream()), "solo")
            torch.cuda.synchronize()
        bad[opt] = int((p != sp).sum().item())
    res = torch.tensor([bad[0], bad[1], bad[2]], device=dev)
    dist.all_reduce(res)
    if rank == 0:
        torch.save({"bad": res.tolist()}, out)
    dist.destroy_process_group()


def test_fused_kernel_bitwise_equals_single_process_reference(tmp_path):
    import torch.multiprocessing as mp
    world = WORLD
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_bitwise_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert torch.load(out)["bad"] == [0, 0, 0]


def test_sync_engine_nvls_matches_p2p(tmp_path):
    """Gradients reduced inside the NVSwitch (`multimem.ld_reduce`) and weights broadcast by it (`multimem.st`): same training
    trajectory as the P2P pull/push kernel up to the bf16 rounding of the reduced gradient; eager and CUDA-graph replay."""
    import torch.multiprocessing as mp
    from distriflow_b200.parallel.symm import multicast_supported
    world = WORLD
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
    world = WORLD
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_async_worker, args=(world, _free_port(), out, graph), nprocs=world, join=True)
    res = torch.load(out)
    for c in res["counters"]:
        # every push to this shard was either applied or refused by the staleness guard: nothing lost
        assert c["applied"] + c["dropped"] == world * res["steps"], c
        # a shard version counts optimizer updates: gradients that met in one accumulation buffer are one update
        assert 1 <= c["version"] <= c["applied"] and c["inflight"] == [0, 0] and c["arrived"] == [0, 0]
        assert c["max_staleness_seen"] <= 4


def _late_rank_worker(rank, world, port, out):
    """Asynchronous mode with a rank that joins late and leaves early (worker churn, abstract_server.ts:108-126): the last
    rank computes nothing until every other rank has FINISHED all its steps, then contributes a few gradients and stops."""
    import time
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_mlp
    from distriflow_b200.parallel.async_engine import AsyncEngine
    torch.manual_seed(0)
    net = mnist_mlp(softmax_output=False).to(dev)
    ce = get_loss("softmaxCrossEntropy")
    eng = AsyncEngine(net, lambda m, x, y: ce(y, m(x)), {"optimizer": "sgd", "learningRate": 1e-3, "maximumStaleness": 64,
                                                          "dtype": "fp32", "gradDtype": "fp32"}, rank=rank, world=world, device=dev)
    late = rank == world - 1
    steps = 3 if late else 10
    flag = out + ".early_done"
    g = torch.Generator().manual_seed(rank)
    waited = 0.0
    if late:
        t0 = time.time()
        while sum(os.path.exists(f"{flag}{r}") for r in range(world - 1)) < world - 1 and time.time() - t0 < 120:
            time.sleep(0.05)
        waited = time.time() - t0
    for _ in range(steps):
        x = torch.rand(32, 28, 28, 1, generator=g).to(dev)
        y = torch.randint(0, 10, (32,), generator=g).to(dev)
        eng.step(x, y)
    torch.cuda.synchronize()                       # the early ranks get here while the late one has not launched a kernel
    if not late:
        open(f"{flag}{rank}", "w").close()
    eng.drain()
    allc = [None] * world
    dist.all_gather_object(allc, {"counters": eng.counters(), "waited": waited, "steps": steps, "lost": eng.state.heaps[0].barrier_failure()})
    if rank == 0:
        torch.save(allc, out)
    dist.destroy_process_group()


def test_async_engine_tolerates_a_rank_that_joins_late_and_leaves_early(tmp_path):
    import torch.multiprocessing as mp
    world = WORLD
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_late_rank_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = torch.load(out)
    assert res[-1]["waited"] < 100, "the early ranks never finished: they were blocked by the absent one"
    pushes = sum(r["steps"] for r in res)
    for r in res:
        c = r["counters"]
        assert c["applied"] + c["dropped"] == pushes, c           # nothing lost, nothing double counted
        assert c["inflight"] == [0, 0] and c["arrived"] == [0, 0]
        assert r["lost"] is None                                  # no barrier gave up: nobody ever waited for the late rank


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
    world = WORLD
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    out = str(tmp_path / "res.pt")
    mp.spawn(_stress_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = torch.load(out)
    assert res["bad"] == 0, res
