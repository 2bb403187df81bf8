"""Collective / async / work-queue kernels on ONE GPU with emulated ranks (one launch per rank over R heaps in the same
process, cross-rank barriers disabled), checked against plain PyTorch math.  Real multi-GPU runs: test_multigpu.py."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

OPT = {"sgd": 0, "momentum": 1, "adam": 2, "none": 3}


def dev():
    return torch.device("cuda", 0)


def _setup(R, n, gdt, pdt):
    from distriflow_b200.parallel.symm import SymmetricHeap
    gsz, psz = torch.empty((), dtype=gdt).element_size(), torch.empty((), dtype=pdt).element_size()
    heaps = SymmetricHeap.emulated(R, n * (gsz + psz) + 8 * n + 65536)
    offs = {}
    for h in heaps:
        offs["g"] = h.alloc(n * gsz)
        offs["p"] = h.alloc(n * psz)
    return heaps, offs


def _ref_opt(kind, w, g, m, v, lr, step, mom=0.9, b1=0.9, b2=0.999, eps=1e-8):
    if kind == "sgd":
        return w - lr * g, m, v
    if kind == "momentum":
        m = mom * m + g
        return w - lr * m, m, v
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    return w - lr * (m / (1 - b1 ** step)) / ((v / (1 - b2 ** step)).sqrt() + eps), m, v


@pytest.mark.parametrize("R", [1, 2, 4, 8])
@pytest.mark.parametrize("kind", ["sgd", "momentum", "adam"])
@pytest.mark.parametrize("gdt,pdt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16)])
def test_sync_rs_opt_ag_emulated(R, kind, gdt, pdt):
    from distriflow_b200.ops._lib import check, dt, ptr, stream
    from distriflow_b200.parallel.engine import _comm_sigs
    L = _comm_sigs()
    torch.manual_seed(R)
    shard = 64 * 37
    n = shard * R
    heaps, offs = _setup(R, n, gdt, pdt)
    grads = [torch.randn(n, device=dev()) for _ in range(R)]
    w0 = torch.randn(n, device=dev())
    for r, h in enumerate(heaps):
        h.view(offs["g"], (n,), gdt).copy_(grads[r])
        h.view(offs["p"], (n,), pdt).copy_(w0)
    gq = [h.view(offs["g"], (n,), gdt).float() for h in heaps]
    gmean = torch.zeros(n, device=dev())
    for q in range(R):
        gmean += gq[q]                                   # same fixed order as the kernel
    gmean = gmean * (1.0 / R)
    masters = [w0[r * shard:(r + 1) * shard].clone() for r in range(R)]
    ms = [torch.zeros(shard, device=dev()) for _ in range(R)]
    vs = [torch.zeros(shard, device=dev()) for _ in range(R)]
    wr, mr, vr = w0.clone(), torch.zeros(n, device=dev()), torch.zeros(n, device=dev())
    for step in (1, 2):
        for r, h in enumerate(heaps):
            check(L.dfb_sync_rs_opt_ag(h.ptrs(offs["g"]), h.ptrs(offs["p"]), None, ptr(masters[r]), ptr(ms[r]), ptr(vs[r]), None,
                                       n, shard, r, R, 0, 1, dt(heaps[0].view(offs["g"], (1,), gdt)),
                                       dt(heaps[0].view(offs["p"], (1,), pdt)), OPT[kind], 0.01, 0.9, 0.9, 0.999, 1e-8, 0.0, step,
                                       8, 0, 0, stream()), "sync")
        torch.cuda.synchronize()
        wr, mr, vr = _ref_opt(kind, wr, gmean, mr, vr, 0.01, step)
        full_master = torch.cat(masters)
        assert torch.allclose(full_master, wr, rtol=2e-5, atol=2e-6)
        for h in heaps:                                  # every rank received every slice (the all-gather half)
            got = h.view(offs["p"], (n,), pdt)
            assert torch.equal(got, full_master.to(pdt))
    for h in heaps:
        h.close()


def test_reduced_out_and_no_push():
    from distriflow_b200.ops._lib import check, dt, ptr, stream
    from distriflow_b200.parallel.engine import _comm_sigs
    L = _comm_sigs()
    R, shard = 4, 256
    n = R * shard
    heaps, offs = _setup(R, n, torch.bfloat16, torch.bfloat16)
    gs = [torch.randn(n, device=dev()).bfloat16() for _ in range(R)]
    for r, h in enumerate(heaps):
        h.view(offs["g"], (n,), torch.bfloat16).copy_(gs[r])
        h.view(offs["p"], (n,), torch.bfloat16).zero_()
    red = torch.empty(shard, device=dev(), dtype=torch.bfloat16)
    master = torch.zeros(shard, device=dev())
    check(L.dfb_sync_rs_opt_ag(heaps[1].ptrs(offs["g"]), heaps[1].ptrs(offs["p"]), None, ptr(master), None, None, ptr(red), n, shard,
                               1, R, 0, 0, 1, 1, OPT["none"], 0.0, 0.0, 0.9, 0.999, 1e-8, 0.0, 1, 4, 0, 0, stream()), "sync")
    torch.cuda.synchronize()
    ref = sum(g.float() for g in gs)[shard:2 * shard] / R
    assert torch.allclose(red.float(), ref, rtol=1e-2, atol=1e-2)
    for h in heaps:
        assert float(h.view(offs["p"], (n,), torch.bfloat16).abs().sum()) == 0.0      # OPT_NONE pushes nothing
        h.close()


def test_allgather_push_emulated():
    from distriflow_b200.ops._lib import check, stream
    from distriflow_b200.parallel.engine import _comm_sigs
    L = _comm_sigs()
    R, shard = 4, 4096
    heaps, offs = _setup(R, R * shard, torch.float32, torch.float32)
    for r, h in enumerate(heaps):
        h.view(offs["p"], (R * shard,), torch.float32).fill_(-1)
        h.view(offs["p"], (R * shard,), torch.float32)[r * shard:(r + 1) * shard] = float(r)
    for r, h in enumerate(heaps):
        check(L.dfb_allgather_push(h.ptrs(offs["p"]), None, R * shard * 4, shard * 4, r, R, 0, 4, stream()), "ag")
    torch.cuda.synchronize()
    want = torch.arange(R, device=dev()).repeat_interleave(shard).float()
    for h in heaps:
        assert torch.equal(h.view(offs["p"], (R * shard,), torch.float32), want)
        h.close()


def test_async_push_apply_pull_invariants():
    """No update lost, staleness never above the bound, version monotone."""
    from distriflow_b200.parallel.async_engine import AsyncShardState
    R, shard = 4, 64 * 8
    n = R * shard
    st = AsyncShardState.emulated(R, n, grad_dtype=torch.float32, param_dtype=torch.float32, lr=0.1, max_staleness=2)
    w0 = torch.randn(n, device=dev())
    st.init_weights(w0)
    ref = w0.clone()
    torch.manual_seed(0)
    for w in range(R):
        st.pull(w)
    applied = 0
    # workers 0..2 push; nobody applies yet -> all three arrive in the same buffer
    gs = [torch.randn(n, device=dev()) for _ in range(R)]
    for w in range(3):
        st.push(w, gs[w])
    for o in range(R):
        st.apply(o)
    applied += 3
    ref -= 0.1 * (gs[0] + gs[1] + gs[2])
    ctl = st.read_ctl(0)
    assert ctl["version"] == 3 and ctl["applied"] == 3 and ctl["inflight"] == [0, 0] and ctl["arrived"] == [0, 0]
    assert torch.allclose(st.full_master(), ref, rtol=1e-5, atol=1e-6)
    # worker 3 still holds version 0 -> staleness 3 > bound 2 -> refused everywhere
    st.push(3, gs[3])
    for o in range(R):
        st.apply(o)
    ctl = st.read_ctl(1)
    assert ctl["version"] == 3 and ctl["dropped"] == 1
    assert torch.allclose(st.full_master(), ref, rtol=1e-5, atol=1e-6)
    # after a pull it is fresh again and goes through; working weights follow the master
    st.pull(3)
    assert torch.allclose(st.local_weights(3), ref, rtol=1e-5, atol=1e-6)
    st.push(3, gs[3])
    for o in range(R):
        st.apply(o)
    ref -= 0.1 * gs[3]
    for o in range(R):
        c = st.read_ctl(o)
        assert c["version"] == 4 and c["applied"] == 4 and c["max_staleness_seen"] <= 2
    assert torch.allclose(st.full_master(), ref, rtol=1e-5, atol=1e-6)
    st.close()


def test_device_work_queue_matches_host_dispatcher():
    from distriflow_b200 import DistriDataset
    from distriflow_b200.parallel.workqueue import DeviceWorkQueue
    batches, epochs = 37, 2
    x = torch.arange(batches * 4, dtype=torch.float32).reshape(batches * 4, 1)
    host = DistriDataset(x, x.clone(), {"batchSize": 4, "epochs": epochs}, use_uuids=False)
    wq = DeviceWorkQueue.local(batches, epochs, device=dev())
    # FCFS order and epoch roll-over identical to the host dispatcher
    for _ in range(batches * epochs):
        hb = host.next()
        t = wq.take()
        assert (t.batch, t.epoch) == (hb.batch, hb.epoch)
        host.complete_batch(hb.batch, hb.epoch)
        wq.complete(t)
    assert host.next().done and wq.take().done
    # stragglers are re-issued and the epoch waits for them
    wq = DeviceWorkQueue.local(5, 1, device=dev())
    got = [wq.take() for _ in range(5)]
    for t in got:
        if t.batch not in (1, 3):
            wq.complete(t)
    again = {wq.take().batch for _ in range(4)}
    assert again == {1, 3}
    wq.complete(got[1]); wq.complete(got[1])           # idempotent
    assert not wq.take().done
    wq.complete(got[3])
    assert wq.take().done


def test_work_queue_stale_duplicate_completion_cannot_poison_the_next_epoch():
    """A duplicate completion of a re-issued batch that arrives after its epoch has rolled over must not set a bit in the
    next epoch's bitmap (the epoch then never completes and every taker spins)."""
    from distriflow_b200.parallel.workqueue import DeviceWorkQueue
    wq = DeviceWorkQueue.local(4, 3, device=dev())
    first = [wq.take() for _ in range(4)]
    dup = wq.take()                                     # fresh tickets are gone: batch 0 is re-issued
    assert dup.batch == 0 and dup.epoch == 0
    for t in first:
        wq.complete(t)                                  # the fourth completion rolls the queue to epoch 1
    wq.complete(dup)                                    # late duplicate from epoch 0: ignored
    wq.complete(first[2])
    second = [wq.take() for _ in range(4)]
    assert [(t.batch, t.epoch) for t in second] == [(0, 1), (1, 1), (2, 1), (3, 1)]
    for t in second:
        wq.complete(t)
    third = [wq.take() for _ in range(4)]
    assert [(t.batch, t.epoch) for t in third] == [(0, 2), (1, 2), (2, 2), (3, 2)]
    for t in third:
        wq.complete(t)
    assert wq.take().done


def test_work_queue_bitmap_is_sized_from_the_dataset():
    """60 000 examples at batch size 1 (a federated-style run over MNIST): the bitmap does not fit the 4 KiB signal pad and
    moves to a heap block."""
    from distriflow_b200.parallel.workqueue import DeviceWorkQueue
    wq = DeviceWorkQueue.local(60000, 1, device=dev())
    t0, t1 = wq.take(), wq.take()
    assert (t0.batch, t1.batch) == (0, 1)
    wq.complete(t0)
    wq.complete(t1)
    assert wq.take().batch == 2 and not wq.take().done


def test_gather_batch_from_ticket():
    from distriflow_b200.parallel.workqueue import DeviceWorkQueue
    data = torch.arange(40 * 8, device=dev(), dtype=torch.float32).reshape(40, 8)
    wq = DeviceWorkQueue.local(10, 1, device=dev())
    wq.take(); wq.take()
    t = wq.take()                                      # batch 2
    out = wq.gather(data, 4)
    assert t.batch == 2 and torch.equal(out, data[8:12])


@pytest.mark.parametrize("R,pad_after", [(2, 0), (4, 0), (8, 0), (4, 5), (8, 3)])
def test_allgather_fused_gemm_emulated(R, pad_after):
    """Each emulated rank holds only its own slice of W; the fused GEMM must pull the rest and produce x @ W.
    `pad_after` > 0: W is a small part of a large bucket (the one-bucket layout), so it lies inside a few shards only and
    some ranks own none of it."""
    from distriflow_b200.ops.dense import _init_sigs, ag_gemm
    from distriflow_b200.parallel.ag_fusion import build_plan
    from distriflow_b200.parallel.symm import SymmetricHeap
    _init_sigs()
    torch.manual_seed(R)
    M, K, N = 1024, 768, 2304
    w_off = 64 * 5                                    # W does not start at the bucket start
    numel = K * N
    total = ((w_off + numel * (1 + pad_after) + 64 * 3 + R * 64 - 1) // (R * 64)) * (R * 64)
    shard = total // R
    heaps = SymmetricHeap.emulated(R, total * 2 + 65536)
    offs = [h.alloc(total * 2) for h in heaps][0]
    W = (torch.randn(K, N, device=dev()) * 0.05).bfloat16()
    x = torch.randn(M, K, device=dev()).bfloat16()
    bias = torch.randn(N, device=dev())
    flat = torch.zeros(total, device=dev(), dtype=torch.bfloat16)
    flat[w_off:w_off + numel] = W.reshape(-1)
    for r, h in enumerate(heaps):                     # every rank: own slice correct, everything else poisoned
        buf = h.view(offs, (total,), torch.bfloat16)
        buf.fill_(float("nan"))
        buf[r * shard:(r + 1) * shard] = flat[r * shard:(r + 1) * shard]
    ref = torch.relu(x.float() @ W.float() + bias)
    for r, h in enumerate(heaps):
        bases = [hh.base + offs for hh in heaps]
        plan = build_plan(K, N, 2, w_off, shard, r, R, bases, dev())
        w_local = h.view(offs, (total,), torch.bfloat16)[w_off:w_off + numel].view(K, N)
        assert plan.remote_bytes > 0 or (r * shard <= w_off and w_off + numel <= (r + 1) * shard)
        y = ag_gemm(x, w_local, plan, bias, "relu")
        torch.cuda.synchronize()
        assert torch.equal(w_local, W), "gathered weights differ"
        err = float((y.float() - ref).abs().max() / ref.abs().max())
        assert err < 1e-2, err
    for h in heaps:
        h.close()


def test_barrier_gives_up_on_a_silent_rank_and_names_it():
    """Failure detection between GPUs: a rank that never launches its kernel must not leave the others spinning for ever.
    Rank 0 of two emulated ranks enters a barrier alone with a 30 ms bound: the kernel ends, the pad names rank 1."""
    import time
    from distriflow_b200.parallel.engine import _comm_sigs
    from distriflow_b200.parallel.symm import SymmetricHeap
    L = _comm_sigs()
    heaps = SymmetricHeap.emulated(2, 65536)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    try:
        heaps[0].set_barrier_timeout(0.03)
        t0 = time.time()
        assert L.dfb_rank_barrier(heaps[0].ptrs(0), 0, 2, 1, st) == 0
        torch.cuda.synchronize()
        assert time.time() - t0 < 5.0
        assert heaps[0].barrier_failure() == (1, 1)
        assert heaps[0].barrier_failure() is None                      # reading clears the record
        # rank 1 arrives later: rank 0's flag of generation 1 is already there, nothing to wait for, nothing recorded
        assert L.dfb_rank_barrier(heaps[1].ptrs(0), 1, 2, 1, st) == 0
        torch.cuda.synchronize()
        assert heaps[1].barrier_failure() is None
    finally:
        heaps[0].set_barrier_timeout(20.0)
        torch.cuda.synchronize()
        for h in heaps:
            h.close()
